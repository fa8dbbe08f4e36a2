"""GPU test of rosy_tex_mips_device_static, the mip-chain entry point that allocates nothing.

Some users of the library cannot get device memory at call time: a plug-in inside an engine or a tensor
runtime that has taken all of HBM into its own arena (cudaMalloc fails there), or a frame / streaming thread
that may not risk the device-wide stall of cudaMalloc / cudaFree.  For them the library carries a fixed 5 MiB
scratch array in its own image and this entry point builds the chain over it.  Checked here, through the C
ABI: it gives the bytes of rosy_tex_mips_device (which takes caller-allocated scratch); it still does when
calls follow one another on different textures, i.e. over the stale contents of the call before; it leaves the
amount of free device memory untouched; and it refuses a texture that does not fit with a clear message.

This file is named to be collected last: `pytest -x` stops at the first failure, and a problem here should not
hide the parity tests.  It passes on its own in a fresh process (the new entry point is then the very first
call into the library, before any rosy_tex_init) as well as after the rest of the suite."""
import numpy as np
import pytest

from conftest import max_abs_diff

pytestmark = pytest.mark.gpu


def _upload(torch, dev, img):
    return torch.from_numpy(img.reshape(-1).copy()).to(dev)


def _static_chain(tex, torch, dev, img, kind):
    h, w = img.shape[:2]
    d = tex.desc(w, h, kind)
    t_in = _upload(torch, dev, img)
    t_chain = torch.full((tex.chain_size(w, h),), 0xA5, dtype=torch.uint8, device=dev)
    tex.mips_device_static(d, t_in.data_ptr(), t_chain.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return t_chain.cpu().numpy()


def _existing_chain(tex, torch, dev, img, kind):
    h, w = img.shape[:2]
    d = tex.desc(w, h, kind)
    t_in = _upload(torch, dev, img)
    t_chain = torch.full((tex.chain_size(w, h),), 0xA5, dtype=torch.uint8, device=dev)
    t_scr = torch.zeros(tex.device_scratch_size(d), dtype=torch.uint8, device=dev)
    tex.mips_device(d, t_in.data_ptr(), t_chain.data_ptr(), t_scr.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return t_chain.cpu().numpy()


def _written(tex, img, kind):
    # colour maps: level 0 is read from the input and not copied into the chain buffer
    return 0 if kind == tex.KIND_NORMAL_MAP else img.size


def test_static_scratch_entry_point_matches_the_allocating_one(tex, oracle, synth):
    import torch

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    # sides that are not multiples of 64: rosy_tex_mips_device runs the per-level kernels as well -> same bytes.
    # A big texture first and a small one of the other kind after it, so that the second call finds the first
    # call's values all over the shared array; then the first again.
    cases = [(synth.albedo(1000, 700, 7), tex.KIND_COLOR),
             (synth.normal(200, 120, 1), tex.KIND_NORMAL_MAP),
             (synth.albedo(1000, 700, 7), tex.KIND_COLOR),
             (synth.make("metal_rough", 333, 77, 2), tex.KIND_METALLIC_ROUGHNESS)]
    # the new entry point runs FIRST (two calls in a row and more), the reference results only afterwards
    got = [_static_chain(tex, torch, dev, img, kind) for img, kind in cases]
    for (img, kind), g in zip(cases, got):
        want = _existing_chain(tex, torch, dev, img, kind)
        first = _written(tex, img, kind)
        assert g.size == want.size
        assert (g[first:] == want[first:]).all(), (img.shape, kind, max_abs_diff(g[first:], want[first:]))
    assert (got[0] == got[2]).all()

    # the largest texture it takes (1024 x 1024 fills the array exactly).  rosy_tex_mips_device uses the fused
    # TMA kernel here: both are held to 1 LSB of the CPU oracle, the tolerance of tests/test_gpu_parity.py
    img = synth.albedo(1024, 1024, 4)
    g = _static_chain(tex, torch, dev, img, tex.KIND_COLOR)
    want = _existing_chain(tex, torch, dev, img, tex.KIND_COLOR)
    ref = np.concatenate([l.reshape(-1) for l in oracle.mip_chain(img, oracle.KIND_SRGB)])
    first = img.size
    assert max_abs_diff(g[first:], ref[first:]) <= 1
    assert max_abs_diff(want[first:], ref[first:]) <= 1
    assert max_abs_diff(g[first:], want[first:]) <= 2

    # a repeat of a call that has already run: free device memory is the same before and after
    img, kind = cases[0]
    h, w = img.shape[:2]
    d = tex.desc(w, h, kind)
    t_in = _upload(torch, dev, img)
    t_chain = torch.zeros(tex.chain_size(w, h), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    launches_before = tex.kernel_launches()
    tex.mips_device_static(d, t_in.data_ptr(), t_chain.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free_before
    assert tex.kernel_launches() - launches_before == tex.count_mipmaps(w, h) - 1  # one kernel per level below 0
    assert (t_chain.cpu().numpy()[img.size:] == got[0][img.size:]).all()


def test_static_scratch_entry_point_refuses_what_does_not_fit(tex, synth):
    import torch

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    w, h = 2048, 1024  # twice the pixels the array is sized for
    t_in = torch.zeros(w * h * 4, dtype=torch.uint8, device=dev)
    t_chain = torch.zeros(tex.chain_size(w, h), dtype=torch.uint8, device=dev)
    with pytest.raises(tex.RosyTexError) as e:
        tex.mips_device_static(tex.desc(w, h, tex.KIND_COLOR), t_in.data_ptr(), t_chain.data_ptr(), 0)
    assert e.value.status == 3  # ROSY_TEX_ERR_ARG
    assert "too large" in str(e.value) and "ROSY_TEX_STATIC_SCRATCH_BYTES" in str(e.value)
    # and the same shape as two halves goes through
    half = synth.albedo(1024, 1024, 0)
    assert _static_chain(tex, torch, dev, half, tex.KIND_COLOR).size == tex.chain_size(1024, 1024)
