"""The CUDA kernel's search code, compiled for the host, against the frozen spec-v2 encoder vectors.

rosy_b200/csrc/bc7_encode.cu builds with g++ under -DBC7_HOST_EMU (tools/bc7_host_emu: intrinsics restated for
the host, one block at a time).  This is a CPU-side check of the KERNEL'S OWN source - prepare / refine / evaluate /
estimate / pack - not of the oracle: a change to the kernel that alters any decision fails here, on a machine
without a GPU, before the -m gpu parity tests see it.  The product never loads this library.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EMU = os.path.join(ROOT, "tools", "bc7_host_emu")


@pytest.fixture(scope="module")
def emu():
    if not os.path.exists("/usr/local/cuda/include/cuda_runtime.h"):
        pytest.skip("CUDA toolkit headers not present")
    subprocess.check_call([os.path.join(EMU, "build.sh")])
    return C.CDLL(os.path.join(EMU, "libbc7emu.so"))


def _encode(emu, img, q):
    h, w = img.shape[:2]
    out = np.zeros((((w + 3) // 4) * ((h + 3) // 4), 16), np.uint8)
    u8p = C.POINTER(C.c_uint8)
    assert emu.bc7_emu_encode_image(img.ctypes.data_as(u8p), w, h, q, out.ctypes.data_as(u8p)) == 0
    return out


def test_kernel_source_reproduces_frozen_spec_v2_on_the_host(emu):
    g = np.load(os.path.join(HERE, "golden", "bc7_spec_v2.npz"))
    px = g["px"]
    n = px.shape[0]
    img = px.reshape(n // 16, 16, 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(n // 16 * 4, 64, 4).copy()
    for name, q in (("fast", 0), ("normal", 1), ("high", 2)):
        bad = np.nonzero((_encode(emu, img, q) != g[name]).any(axis=1))[0]
        assert bad.size == 0, (name, bad[:8])


def test_kernel_source_matches_oracle_on_ragged_levels(emu, oracle, synth):
    for name, idx, (w, h) in [("albedo", 15, (33, 31)), ("random", 9, (20, 12)), ("normal", 1, (64, 64)), ("albedo", 7, (7, 3))]:
        for lvl in oracle.mip_chain(synth.make(name, w, h, idx), oracle.KIND_SRGB)[:3]:
            lvl = np.ascontiguousarray(lvl)
            for q in (0, 1, 2):
                want = oracle.encode_image(lvl, oracle.FMT_BC7, q).reshape(-1, 16)
                assert (_encode(emu, lvl, q) == want).all(), (name, lvl.shape, q)
