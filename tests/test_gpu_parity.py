"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle.

Bars (north_star / SURVEY.md 8c):
  * mip levels: identical extents and count, <= 1 LSB per 8-bit channel
  * BC4 / BC5: bit exact (integer algorithm of stb_dxt.h)
  * BC7: every block valid and decodable; PSNR >= oracle - 0.1 dB at the matched tier.  The
    kernel implements the same search specification as the oracle, so we additionally require
    bit-exact blocks and report it separately from the PSNR gate.
"""
import os

import numpy as np
import pytest

from conftest import max_abs_diff

pytestmark = pytest.mark.gpu

KINDS = [("albedo", 0), ("albedo", 7), ("normal", 1), ("metal_rough", 2), ("random", 3)]


def _chain_check(tex, oracle, img, kind, normalize=-1):
    got = tex.mips_rgba8(img, kind, normalize)
    want = oracle.mip_chain(img, oracle.KIND_NORMAL if kind == tex.KIND_NORMAL_MAP else oracle.KIND_SRGB,
                            oracle.NORMALIZE_UNPACK if normalize != 0 else oracle.NORMALIZE_OFF)
    assert len(got) == len(want) == oracle.count_mipmaps(img.shape[1], img.shape[0])
    worst = 0
    for l, (g, w) in enumerate(zip(got, want)):
        assert g.shape == w.shape, f"level {l} extent {g.shape} != {w.shape}"
        d = max_abs_diff(g, w)
        assert d <= 1, f"level {l}: max abs diff {d} LSB"
        worst = max(worst, d)
    return worst


@pytest.mark.parametrize("size", [(256, 256), (1024, 1024), (512, 128), (64, 64)])
@pytest.mark.parametrize("name,index", [("albedo", 0), ("albedo", 7), ("metal_rough", 2), ("random", 3)])
def test_colour_mip_chain_within_1lsb(tex, oracle, synth, size, name, index):
    img = synth.make(name, size[0], size[1], index)
    _chain_check(tex, oracle, img, tex.KIND_COLOR)


@pytest.mark.parametrize("size", [(256, 256), (1024, 1024), (128, 512)])
@pytest.mark.parametrize("normalize", [1, 0])
def test_normal_mip_chain_within_1lsb(tex, oracle, synth, size, normalize):
    img = synth.normal(size[0], size[1], 1)
    _chain_check(tex, oracle, img, tex.KIND_NORMAL_MAP, normalize)


@pytest.mark.parametrize("size", [(1, 1), (2, 2), (3, 5), (17, 9), (100, 60), (255, 257), (640, 360), (1, 64), (64, 1)])
def test_mip_chain_ragged_extents(tex, oracle, synth, size):
    """odd / non-square / non-power-of-two: generic path, polyphase box (H7)"""
    w, h = size
    _chain_check(tex, oracle, synth.albedo(w, h, 7 if w > 8 else 0), tex.KIND_COLOR)
    _chain_check(tex, oracle, synth.normal(w, h, 1), tex.KIND_NORMAL_MAP)


def test_fused_and_generic_mip_paths_agree(tex, oracle, synth, monkeypatch):
    """the TMA-fused single-launch kernel and the per-level kernels are both within 1 LSB"""
    import subprocess, sys, os, textwrap

    code = textwrap.dedent(
        """
        import numpy as np, sys
        sys.path.insert(0, %r)
        from rosy_b200 import texture as T, synth as S
        import oracle as O
        img = S.albedo(512, 512, 7)
        got = T.mips_rgba8(img, T.KIND_COLOR)
        want = O.mip_chain(img, O.KIND_SRGB)
        print(max(int(np.abs(g.astype(int) - w.astype(int)).max()) for g, w in zip(got, want)))
        """
        % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    )
    for mode in ("generic", "fused"):
        env = dict(os.environ, ROSY_TEX_MIPS=mode)
        out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr[-2000:]
        assert int(out.stdout.strip().splitlines()[-1]) <= 1, (mode, out.stdout)


def test_known_answers(tex, synth):
    """C6: constants survive every level; flat normal stays (128,128,255)"""
    levels = tex.mips_rgba8(synth.constant(256, 256, (187, 187, 187, 255)), tex.KIND_COLOR)
    for l in levels:
        assert (l == np.array([187, 187, 187, 255], dtype=np.uint8)).all()
    levels = tex.mips_rgba8(synth.constant(128, 128, (128, 128, 255, 255)), tex.KIND_NORMAL_MAP)
    for l in levels:
        assert (l == np.array([128, 128, 255, 255], dtype=np.uint8)).all()
    # fully transparent texels carry no colour into the mips; an all-transparent image gives rgb = 0
    img = synth.constant(64, 64, (200, 10, 10, 0))
    levels = tex.mips_rgba8(img, tex.KIND_COLOR)
    assert (levels[1] == 0).all()


@pytest.mark.parametrize("fmt_name", ["bc5", "bc4"])
@pytest.mark.parametrize("name,index,size", [("normal", 1, (256, 256)), ("random", 3, (128, 64)), ("albedo", 0, (30, 18))])
def test_bc45_bit_exact(tex, oracle, synth, fmt_name, name, index, size):
    fmt_t = tex.FORMAT_BC5 if fmt_name == "bc5" else tex.FORMAT_BC4
    fmt_o = oracle.FMT_BC5 if fmt_name == "bc5" else oracle.FMT_BC4
    img = synth.make(name, size[0], size[1], index)
    got = tex.compress_level(img, fmt_t)
    want = oracle.encode_image(img, fmt_o)
    assert got.shape == want.shape
    assert (got == want).all()


@pytest.mark.parametrize("quality", [0, 1, 2])
@pytest.mark.parametrize("name,index", KINDS)
def test_bc7_level_parity(tex, oracle, synth, quality, name, index):
    w, h = (128, 128) if quality == 2 else (256, 256)
    img = synth.make(name, w, h, index)
    got = tex.compress_level(img, tex.FORMAT_BC7, quality)
    want = oracle.encode_image(img, oracle.FMT_BC7, quality)
    modes = oracle.bc7_block_modes(got)
    assert (modes >= 0).all(), "reserved BC7 mode emitted"
    dec_g = oracle.decode_image(got, w, h)
    dec_o = oracle.decode_image(want, w, h)
    p_g, p_o = oracle.psnr(img, dec_g), oracle.psnr(img, dec_o)
    assert p_g >= p_o - 0.1, f"PSNR {p_g:.3f} dB < oracle {p_o:.3f} dB - 0.1"
    same = (got.reshape(-1, 16) == want.reshape(-1, 16)).all(axis=1)
    assert same.all(), f"{(~same).sum()} of {same.size} blocks differ from the oracle (PSNR {p_g:.3f} vs {p_o:.3f})"


def test_bc7_partial_blocks_and_tiny_levels(tex, oracle, synth):
    for (w, h) in [(1, 1), (2, 2), (3, 7), (10, 6), (33, 31)]:
        img = synth.albedo(w, h, 7 if w > 4 else 0)
        got = tex.compress_level(img, tex.FORMAT_BC7, 1)
        want = oracle.encode_image(img, oracle.FMT_BC7, 1)
        assert (got == want).all(), (w, h)


def test_bc7_solid_and_two_colour_blocks_exact(tex, oracle):
    rng = np.random.default_rng(5)
    img = np.zeros((64, 64, 4), dtype=np.uint8)
    for by in range(16):
        for bx in range(16):
            img[4 * by : 4 * by + 4, 4 * bx : 4 * bx + 4] = rng.integers(0, 256, 4)
    got = tex.compress_level(img, tex.FORMAT_BC7, 1)
    dec = oracle.decode_image(got, 64, 64)
    assert (dec == img).all(), "solid blocks must decode exactly"
    # opaque two-colour blocks: both colours are end points of some mode -> exact
    img2 = np.zeros((4, 64, 4), dtype=np.uint8)
    for bx in range(16):
        a = np.append(rng.integers(0, 128, 3) * 2 + 1, 255)  # odd values: 7-bit + p-bit 1 exact
        b = np.append(rng.integers(0, 128, 3) * 2 + 1, 255)
        blk = np.where(rng.integers(0, 2, (4, 4, 1)) == 1, a, b)
        img2[:, 4 * bx : 4 * bx + 4] = blk
    dec2 = oracle.decode_image(tex.compress_level(img2, tex.FORMAT_BC7, 1), 64, 4)
    assert (dec2 == img2).all()


@pytest.mark.parametrize("kind_name,index,kind,fmt", [("albedo", 0, 0, 0), ("albedo", 7, 0, 0), ("normal", 1, 1, 0),
                                                       ("normal", 1, 1, 1), ("metal_rough", 2, 2, 0), ("metal_rough", 2, 2, 2)])
def test_whole_stage_dds(tex, oracle, synth, kind_name, index, kind, fmt):
    """generate_*_texture memory to memory: container, level sizes, validity, PSNR vs the oracle"""
    w = h = 256
    img = synth.make(kind_name, w, h, index)
    got = tex.encode_rgba8(img, kind, fmt, 1)
    okind = oracle.KIND_NORMAL if kind == 1 else oracle.KIND_SRGB
    want = oracle.generate_texture(img, okind, fmt, 1)
    assert got.size == want.size == oracle.dds_size(w, h, fmt)
    assert (got[:148] == want[:148]).all(), "DDS header differs from the oracle's"
    chain = oracle.mip_chain(img, okind)
    off = 148
    bb = 8 if fmt == 2 else 16
    for l, ref in enumerate(chain):
        lh, lw = ref.shape[:2]
        n = ((lw + 3) // 4) * ((lh + 3) // 4) * bb
        g = got[off : off + n]
        o = want[off : off + n]
        if fmt == 0:
            assert (oracle.bc7_block_modes(g) >= 0).all()
        dg = oracle.decode_image(g, lw, lh, fmt)
        do = oracle.decode_image(o, lw, lh, fmt)
        ch = 4 if fmt == 0 else (2 if fmt == 1 else 1)
        assert oracle.psnr(ref, dg, ch) >= oracle.psnr(ref, do, ch) - 0.1, f"level {l}"
        off += n
    assert off == got.size
    if oracle.have_ref_dds():
        info = oracle.ref_dds_read(got.tobytes())
        assert info["numMips"] == len(chain) and info["width"] == w and info["height"] == h
        assert info["format"] == {0: 98, 1: 83, 2: 80}[fmt]


def test_file_entry_points(tex, oracle, synth, tmp_path):
    """Packager.h:8-9 semantics: <image>.dds appears next to the image; failures raise"""
    from PIL import Image

    img = synth.albedo(128, 64, 7)
    p = tmp_path / "albedo.png"
    Image.fromarray(img).save(p)
    tex.generate_srgb_texture(str(p))
    dds = np.fromfile(tmp_path / "albedo.dds", dtype=np.uint8)
    assert (dds == tex.encode_rgba8(img, tex.KIND_COLOR)).all()
    nrm = synth.normal(64, 64, 1)
    p2 = tmp_path / "n.tga"
    hdr = bytes([0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 64, 0, 64, 0, 32, 0x28])
    p2.write_bytes(hdr + nrm[..., [2, 1, 0, 3]].tobytes())
    tex.generate_normal_map_texture(str(p2))
    dds2 = np.fromfile(tmp_path / "n.dds", dtype=np.uint8)
    assert (dds2 == tex.encode_rgba8(nrm, tex.KIND_NORMAL_MAP)).all()
    from rosy_b200 import RosyTexError

    with pytest.raises(RosyTexError):
        tex.generate_srgb_texture(str(tmp_path / "missing.png"))
    assert not (tmp_path / "missing.dds").exists()


def test_work_queue_matches_sync_and_is_deterministic(tex, synth):
    imgs = [(synth.albedo(256, 256, i), tex.KIND_COLOR) for i in range(6)] + [(synth.normal(128, 128, 9), tex.KIND_NORMAL_MAP)]
    q = tex.Pending()
    for img, kind in imgs:
        q.submit(img, kind)
    outs = q.wait_all()
    for (img, kind), out in zip(imgs, outs):
        assert (out == tex.encode_rgba8(img, kind)).all()


def test_work_queue_batches_small_textures_identically(tex, synth):
    """A backlog of small textures goes through the queue several per launch (worker_main /
    run_host_batch): mixed extents (ragged, 1x1, non power of two), kinds, formats and quality
    tiers must come back with exactly the bytes of the synchronous single-texture call, twice."""
    rng = np.random.RandomState(11)
    sizes = [(256, 256), (512, 512), (100, 60), (4, 4), (1, 1), (7, 300), (1024, 1024), (64, 192), (333, 129), (16, 16)]
    cases = []
    for i in range(90):
        w, h = sizes[rng.randint(len(sizes))]
        k = rng.randint(3)
        if k == 1:
            img, kind = synth.normal(w, h, i), tex.KIND_NORMAL_MAP
            fmt = [tex.FORMAT_BC7, tex.FORMAT_BC5][rng.randint(2)]
        elif k == 2:
            img, kind = synth.make("metal_rough", w, h, i), tex.KIND_METALLIC_ROUGHNESS
            fmt = [tex.FORMAT_BC7, tex.FORMAT_BC4][rng.randint(2)]
        else:
            img, kind, fmt = synth.albedo(w, h, i), tex.KIND_COLOR, tex.FORMAT_BC7
        cases.append((img, kind, fmt, int(rng.randint(3))))
    cases.insert(40, (synth.albedo(2048, 1024, 5), tex.KIND_COLOR, tex.FORMAT_BC7, 1))  # too large for a batch: alone
    want = [tex.encode_rgba8(img, kind, fmt, q) for img, kind, fmt, q in cases]
    for _ in range(2):
        q = tex.Pending()
        for img, kind, fmt, qu in cases:
            q.submit(img, kind, fmt, qu)
        outs = q.wait_all()
        assert len(outs) == len(want)
        for i, (o, w_) in enumerate(zip(outs, want)):
            assert o.size == w_.size and (o == w_).all(), f"texture {i} {cases[i][0].shape} differs from the synchronous call"


def test_work_queue_batch_member_error_is_isolated(tex, synth):
    """one member of a batch with an output buffer that is too small: wait_all reports it, the others are complete"""
    import ctypes as C

    L = tex.lib()
    imgs = [synth.albedo(128, 128, i) for i in range(12)]
    descs = [tex.desc(128, 128, tex.KIND_COLOR, tex.FORMAT_BC7, 1, -1) for _ in imgs]
    outs = [np.zeros(tex.dds_size(d), np.uint8) for d in descs]
    lens = [C.c_size_t() for _ in imgs]
    for i, (img, d, o, n) in enumerate(zip(imgs, descs, outs, lens)):
        cap = 200 if i == 5 else o.size
        assert L.rosy_tex_submit_rgba8(C.byref(d), img.ctypes.data, o.ctypes.data, cap, C.byref(n)) == 0
    assert L.rosy_tex_wait_all() != 0
    assert L.rosy_tex_wait_all() == 0
    for i, (img, o) in enumerate(zip(imgs, outs)):
        if i != 5:
            assert (o == tex.encode_rgba8(img, tex.KIND_COLOR, tex.FORMAT_BC7, 1)).all(), i


def _full_size_parity(tex, oracle, img, kind, fmt, quality, got=None, per_level_min_px=1024):
    """Whole .dds at a BASELINE.json size, block for block.

    (1) exact: every level's blocks == the oracle's encoder run on THE SAME level pixels (the GPU's own mip
        chain, which the mip tests bound to <= 1 LSB of the oracle's) -> the encoder is bit-exact at full size;
    (2) end to end against oracle.generate_texture (the oracle's own chain): blocks may differ only where a
        mip texel differs by its 1 LSB; they are counted, and the per-level PSNR (vs the oracle's uncompressed
        level) must stay within 0.1 dB of the oracle's."""
    h, w = img.shape[:2]
    okind = oracle.KIND_NORMAL if kind == tex.KIND_NORMAL_MAP else oracle.KIND_SRGB
    if got is None:  # else: the .dds bytes some other route produced (the Packager CLI), checked by the same rules
        got = tex.encode_rgba8(img, kind, fmt, quality)
    want = oracle.generate_texture(img, okind, fmt, quality)
    assert got.size == want.size == oracle.dds_size(w, h, fmt) and (got[:148] == want[:148]).all()
    chain_g = tex.mips_rgba8(img, kind)
    chain_o = oracle.mip_chain(img, okind)
    bb = 8 if fmt == oracle.FMT_BC4 else 16
    ch = 4 if fmt == oracle.FMT_BC7 else (2 if fmt == oracle.FMT_BC5 else 1)
    off, differing, total = 148, 0, 0
    sse_g = sse_o = npx = 0
    for l, (lg, lo) in enumerate(zip(chain_g, chain_o)):
        lh, lw = lo.shape[:2]
        n = ((lw + 3) // 4) * ((lh + 3) // 4) * bb
        g, o = got[off : off + n], want[off : off + n]
        exact = oracle.encode_image(lg, fmt, quality)
        bad = np.nonzero((g.reshape(-1, bb) != exact.reshape(-1, bb)).any(axis=1))[0]
        assert bad.size == 0, f"level {l}: {bad.size} blocks differ from the oracle on identical pixels, first {bad[:4]}"
        if fmt == oracle.FMT_BC7:
            assert (oracle.bc7_block_modes(g) >= 0).all()
        differing += int((g.reshape(-1, bb) != o.reshape(-1, bb)).any(axis=1).sum())
        total += n // bb
        assert max_abs_diff(lg, lo) <= 1, f"level {l}"
        # PSNR delta: each encoder against the uncompressed level it was given (they differ by
        # <= 1 LSB); per level where a level has enough texels for a 0.1 dB bar to mean something, and
        # chain-weighted (squared error pooled over every texel of every level) below
        dg = lg[..., :ch].astype(np.int64) - oracle.decode_image(g, lw, lh, fmt)[..., :ch].astype(np.int64)
        do = lo[..., :ch].astype(np.int64) - oracle.decode_image(o, lw, lh, fmt)[..., :ch].astype(np.int64)
        sg, so = int((dg * dg).sum()), int((do * do).sum())
        sse_g += sg
        sse_o += so
        npx += dg.size
        if lw * lh >= per_level_min_px:
            pg = 99.0 if sg == 0 else 10 * np.log10(255.0**2 * dg.size / sg)
            po = 99.0 if so == 0 else 10 * np.log10(255.0**2 * do.size / so)
            assert pg >= po - 0.1, f"level {l}: PSNR {pg:.3f} dB vs oracle {po:.3f} dB"
        off += n
    assert off == got.size
    pg = 99.0 if sse_g == 0 else 10 * np.log10(255.0**2 * npx / sse_g)
    po = 99.0 if sse_o == 0 else 10 * np.log10(255.0**2 * npx / sse_o)
    assert pg >= po - 0.1, f"chain-weighted PSNR {pg:.3f} dB vs oracle {po:.3f} dB"
    # blocks that differ end to end come from 1-LSB mip differences only; they stay a small minority
    assert differing / total < 0.05, f"{differing} of {total} blocks differ end to end"
    return differing, total


def test_config0_1024_albedo_bc7_block_for_block(tex, oracle, synth):
    """BASELINE.json configs[0]: one 1024^2 sRGB albedo, full chain + BC7 (Quality_Normal), whole .dds"""
    d, t = _full_size_parity(tex, oracle, synth.albedo(1024, 1024, 0), tex.KIND_COLOR, tex.FORMAT_BC7, 1)
    print(f"configs[0]: {d} of {t} blocks differ end to end (1-LSB mip differences)")


@pytest.mark.parametrize("fmt", [1, 0])
def test_config1_4096_normal_block_for_block(tex, oracle, synth, fmt):
    """BASELINE.json configs[1]: one 4096^2 normal map, renormalised chain + BC5 (north_star) and BC7 (reference)"""
    d, t = _full_size_parity(tex, oracle, synth.normal(4096, 4096, 1), tex.KIND_NORMAL_MAP, fmt, 1)
    print(f"configs[1] fmt {fmt}: {d} of {t} blocks differ end to end")


@pytest.mark.parametrize("quality", [0, 1, 2])
def test_config4_2048_albedo_each_tier_block_for_block(tex, oracle, synth, quality):
    """BASELINE.json configs[4]: 2048^2 albedo at fast / normal / high (cut-out alpha at normal)"""
    _full_size_parity(tex, oracle, synth.albedo(2048, 2048, 7 if quality == 1 else 3), tex.KIND_COLOR, tex.FORMAT_BC7, quality)


def test_bc7_kernel_reproduces_frozen_spec_v2(tex):
    """the kernel against the committed spec-v2 encoder vectors directly (no oracle in the loop)"""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bc7_spec_v2.npz"))
    px = g["px"]
    n = px.shape[0]
    img = px.reshape(n // 16, 16, 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(n // 16 * 4, 64, 4).copy()
    for name, q in (("fast", 0), ("normal", 1), ("high", 2)):
        got = tex.compress_level(img, tex.FORMAT_BC7, q).reshape(-1, 16)
        bad = np.nonzero((got != g[name]).any(axis=1))[0]
        assert bad.size == 0, (name, bad[:8])


def _level_blocks(dds, w, h, bs, levels):
    """-> per level (blocks_y, blocks_x, bs) views of the payload of a .dds"""
    out, off = [], 148
    for k in range(levels):
        lw, lh = max(1, w >> k), max(1, h >> k)
        bx, by = (lw + 3) // 4, (lh + 3) // 4
        out.append(dds[off : off + bx * by * bs].reshape(by, bx, bs))
        off += bx * by * bs
    return out


@pytest.mark.parametrize("size", [(8192, 8192), (16384, 4096), (4096, 16384)])
@pytest.mark.parametrize("case", ["bc7_colour", "bc5_normal", "bc4_data"])
def test_maximum_sizes_by_periodicity(tex, synth, size, case):
    """Sizes the oracle cannot reach in seconds (config 4's 8192^2 and beyond): a texture made of
    512x512 repeats of a base image has mips and blocks that are the same repeats of the base
    image's (box filter and block search are local, 512 is a multiple of every footprint), for
    levels 0..7 (base extent >= 4).  The base image's output is checked against the oracle by the
    tests above; this one checks addressing and tiling over > 2^26 pixels."""
    w, h = size
    if case == "bc7_colour":
        base, kind, fmt, bs, q = synth.albedo(512, 512, 7), tex.KIND_COLOR, tex.FORMAT_BC7, 16, 0
    elif case == "bc5_normal":
        base, kind, fmt, bs, q = synth.normal(512, 512, 1), tex.KIND_NORMAL_MAP, tex.FORMAT_BC5, 16, 1
    else:
        base, kind, fmt, bs, q = synth.make("metal_rough", 512, 512, 2), tex.KIND_METALLIC_ROUGHNESS, tex.FORMAT_BC4, 8, 1
    big = np.ascontiguousarray(np.tile(base, (h // 512, w // 512, 1)))
    small = _level_blocks(tex.encode_rgba8(base, kind, fmt, q), 512, 512, bs, 8)
    got = tex.encode_rgba8(big, kind, fmt, q)
    levels = tex.count_mipmaps(w, h)
    assert got.size == tex.dds_size(tex.desc(w, h, kind, fmt, q, -1))
    large = _level_blocks(got, w, h, bs, levels)
    for k in range(8):
        want = np.tile(small[k], (h // 512, w // 512, 1))
        assert large[k].shape == want.shape
        assert (large[k] == want).all(), f"level {k} is not the periodic repeat of the base image's level"
    # below the period: every level of a periodic image stays periodic down to 1 texel per period
    chain = tex.mips_rgba8(big, kind)
    base_chain = tex.mips_rgba8(base, kind)
    for k in range(10):
        assert (chain[k] == np.tile(base_chain[k], (h // 512, w // 512, 1))).all(), f"mip level {k}"


@pytest.mark.parametrize("size", [(2048, 2048), (4096, 64), (64, 4096), (192, 320), (4096, 4096), (2048, 1024)])
def test_fused_mip_kernel_shapes(tex, oracle, synth, size):
    """TMA-fused kernel: tails that go through global scratch (4096^2), 1-tile-wide chains, odd tile
    counts (3x5 tiles -> polyphase box in the tail), for colour (opaque and cut-out alpha) and normals"""
    w, h = size
    for name, index, kind, normalize in [("albedo", 0, tex.KIND_COLOR, -1), ("albedo", 7, tex.KIND_COLOR, -1),
                                         ("normal", 1, tex.KIND_NORMAL_MAP, 1), ("normal", 1, tex.KIND_NORMAL_MAP, 0)]:
        if w * h > 4096 * 2048 and (index == 7 or normalize == 0):
            continue  # keep the oracle's share of the run short
        _chain_check(tex, oracle, synth.make(name, w, h, index), kind, normalize)


@pytest.mark.parametrize("size", [(64, 64), (192, 320), (1024, 512), (2048, 2048)])
def test_fused_mip_encode_equals_separate_stages(tex, oracle, synth, size):
    """BC5 / BC4: levels 0..4 are encoded inside the fused mip kernel (one block per thread, per four lanes, and by the
    auxiliary warp); the .dds must hold exactly the blocks the stand-alone encoder gives for the chain that
    rosy_tex_mips_rgba8 returns (same mip arithmetic, stb_dxt block for block), for every image kind"""
    w, h = size
    cases = [("normal", 1, tex.KIND_NORMAL_MAP, tex.FORMAT_BC5, 1), ("normal", 4, tex.KIND_NORMAL_MAP, tex.FORMAT_BC5, 0),
             ("normal", 2, tex.KIND_NORMAL_MAP, tex.FORMAT_BC4, 1), ("metal_rough", 2, tex.KIND_METALLIC_ROUGHNESS, tex.FORMAT_BC4, -1),
             ("albedo", 7, tex.KIND_COLOR, tex.FORMAT_BC5, -1), ("random", 3, tex.KIND_COLOR, tex.FORMAT_BC4, -1)]
    for name, index, kind, fmt, normalize in cases:
        if w * h > 1024 * 1024 and name in ("random", "albedo"):
            continue
        img = synth.make(name, w, h, index)
        got = tex.encode_rgba8(img, kind, fmt, 1, normalize=normalize)
        levels = tex.mips_rgba8(img, kind, normalize=normalize)
        off, bb = 148, (8 if fmt == tex.FORMAT_BC4 else 16)
        for l, lv in enumerate(levels):
            lh, lw = lv.shape[:2]
            n = ((lw + 3) // 4) * ((lh + 3) // 4) * bb
            want = oracle.encode_image(lv, fmt)
            bad = np.nonzero((got[off : off + n].reshape(-1, bb) != want.reshape(-1, bb)).any(axis=1))[0]
            assert bad.size == 0, f"{name} {w}x{h} fmt {fmt} level {l}: {bad.size} blocks differ, first {bad[:4]}"
            off += n
        assert off == got.size


def test_mip_chain_fraction_of_exact_matches(tex, oracle, synth):
    """the <= 1 LSB bar is met with almost every texel identical to the oracle's"""
    for name, index, kind in [("albedo", 0, tex.KIND_COLOR), ("albedo", 7, tex.KIND_COLOR), ("normal", 1, tex.KIND_NORMAL_MAP)]:
        img = synth.make(name, 1024, 1024, index)
        got = tex.mips_rgba8(img, kind)
        want = oracle.mip_chain(img, oracle.KIND_NORMAL if kind == tex.KIND_NORMAL_MAP else oracle.KIND_SRGB)
        diff = sum(int((g != w).sum()) for g, w in zip(got[1:], want[1:]))
        total = sum(g.size for g in got[1:])
        assert diff / total < 0.02, f"{name}: {diff} of {total} channel values differ from the oracle"


def test_batched_device_entry_matches_per_texture(tex, synth):
    """rosy_tex_encode_batch_device: one mip launch for mixed sizes / kinds (plus a ragged texture that
    takes the per-level kernels) gives the bytes of the per-texture entry points"""
    import torch

    dev = torch.device("cuda", 0)
    specs = [("albedo", 0, 256, 256, tex.KIND_COLOR), ("albedo", 7, 512, 128, tex.KIND_COLOR),
             ("normal", 1, 128, 128, tex.KIND_NORMAL_MAP), ("albedo", 2, 100, 60, tex.KIND_COLOR),
             ("metal_rough", 2, 64, 64, tex.KIND_METALLIC_ROUGHNESS), ("normal", 3, 256, 64, tex.KIND_NORMAL_MAP),
             ("albedo", 4, 1024, 1024, tex.KIND_COLOR)]
    keep, entries, want = [], [], []
    for name, index, w, h, kind in specs:
        img = synth.make(name, w, h, index)
        d = tex.desc(w, h, kind, tex.FORMAT_BC7, tex.QUALITY_FAST, -1)
        t_in = torch.from_numpy(img.reshape(-1).copy()).to(dev)
        t_chain = torch.zeros(tex.chain_size(w, h), dtype=torch.uint8, device=dev)
        t_scr = torch.zeros(tex.device_scratch_size(d), dtype=torch.uint8, device=dev)
        t_blk = torch.zeros(tex.dds_size(d) - tex.DDS_HEADER_BYTES, dtype=torch.uint8, device=dev)
        keep.append((t_in, t_chain, t_scr, t_blk))
        entries.append((d, t_in.data_ptr(), t_chain.data_ptr(), t_scr.data_ptr(), t_blk.data_ptr()))
        want.append((tex.mips_rgba8(img, kind), tex.encode_rgba8(img, kind, tex.FORMAT_BC7, tex.QUALITY_FAST), kind, img))
    batch = tex.make_batch(entries)
    t_batch = torch.zeros(tex.batch_scratch_size(len(entries)), dtype=torch.uint8, device=dev)
    for _ in range(2):  # twice: the tickets must return to zero
        tex.encode_batch_device(batch, t_batch.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
    for (t_in, t_chain, t_scr, t_blk), (levels, dds, kind, img) in zip(keep, want):
        chain = t_chain.cpu().numpy()
        ref = np.concatenate([l.reshape(-1) for l in levels])
        first = 0 if kind == tex.KIND_NORMAL_MAP else img.size  # colour level 0 is read from the input
        assert (chain[first:] == ref[first:]).all()
        assert (t_blk.cpu().numpy() == dds[tex.DDS_HEADER_BYTES:]).all()


@pytest.mark.gpu
def test_batched_blocks_mixed_tiers_and_formats(tex, synth):
    """the batched BC7 launch groups textures by quality tier (one launch each) and leaves BC5 / BC4
    textures to their own launches; CTAs that straddle two textures must not mix them up"""
    import torch

    dev = torch.device("cuda", 0)
    specs = [("albedo", 0, 64, 64, tex.KIND_COLOR, tex.FORMAT_BC7, tex.QUALITY_NORMAL),
             ("albedo", 5, 100, 60, tex.KIND_COLOR, tex.FORMAT_BC7, tex.QUALITY_FAST),
             ("normal", 1, 128, 64, tex.KIND_NORMAL_MAP, tex.FORMAT_BC5, tex.QUALITY_NORMAL),
             ("albedo", 7, 256, 128, tex.KIND_COLOR, tex.FORMAT_BC7, tex.QUALITY_NORMAL),
             ("metal_rough", 2, 64, 192, tex.KIND_METALLIC_ROUGHNESS, tex.FORMAT_BC4, tex.QUALITY_NORMAL),
             ("random", 3, 68, 68, tex.KIND_COLOR, tex.FORMAT_BC7, tex.QUALITY_HIGH),
             ("normal", 4, 64, 64, tex.KIND_NORMAL_MAP, tex.FORMAT_BC7, tex.QUALITY_FAST)]
    keep, entries, want = [], [], []
    for name, index, w, h, kind, fmt, q in specs:
        img = synth.make(name, w, h, index)
        d = tex.desc(w, h, kind, fmt, q, -1)
        t_in = torch.from_numpy(img.reshape(-1).copy()).to(dev)
        t_chain = torch.zeros(tex.chain_size(w, h), dtype=torch.uint8, device=dev)
        t_scr = torch.zeros(tex.device_scratch_size(d), dtype=torch.uint8, device=dev)
        t_blk = torch.zeros(tex.dds_size(d) - tex.DDS_HEADER_BYTES, dtype=torch.uint8, device=dev)
        keep.append((t_in, t_chain, t_scr, t_blk))
        entries.append((d, t_in.data_ptr(), t_chain.data_ptr(), t_scr.data_ptr(), t_blk.data_ptr()))
        want.append(tex.encode_rgba8(img, kind, fmt, q))
    batch = tex.make_batch(entries)
    t_batch = torch.zeros(tex.batch_scratch_size(len(entries)), dtype=torch.uint8, device=dev)
    tex.encode_batch_device(batch, t_batch.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for (t_in, t_chain, t_scr, t_blk), dds in zip(keep, want):
        assert (t_blk.cpu().numpy() == dds[tex.DDS_HEADER_BYTES:]).all()


@pytest.mark.gpu
def test_file_queue_many_files_and_one_bad(tex, synth, tmp_path):
    """rosy_tex_submit: decode threads -> pinned staging -> device slots -> writer threads.  Every good file
    gets the bytes of the synchronous entry point; a corrupt file is reported by wait_all and leaves no .dds"""
    from PIL import Image

    from rosy_b200 import RosyTexError

    paths = []
    for i in range(10):
        img = synth.make(["albedo", "normal", "metal_rough"][i % 3], 64 + 64 * (i % 3), 128, i)
        p = tmp_path / (f"t{i}.jpg" if i % 4 == 1 else f"t{i}.png")
        (Image.fromarray(img[..., :3]) if p.suffix == ".jpg" else Image.fromarray(img)).save(p)
        paths.append((p, tex.KIND_NORMAL_MAP if i % 3 == 1 else tex.KIND_COLOR))
    bad = tmp_path / "bad.png"
    bad.write_bytes(b"\x89PNG\r\n\x1a\n" + b"junk" * 8)
    for k, (p, kind) in enumerate(paths):
        if k == 5:
            tex.submit_file(str(bad), tex.KIND_COLOR)
        tex.submit_file(str(p), kind)
    with pytest.raises(RosyTexError) as e:
        tex.wait_all()
    assert e.value.status == 2
    assert not (tmp_path / "bad.dds").exists()
    for p, kind in paths:
        want = tex.encode_rgba8(tex.load_image(str(p)), kind)
        got = np.fromfile(p.with_suffix(".dds"), dtype=np.uint8)
        assert got.size == want.size and (got == want).all(), p.name
    tex.wait_all()  # the error was consumed


@pytest.mark.gpu
def test_opt_in_linear_data_maps_and_srgb_flag(tex, oracle, synth):
    """N4 (off by default): metallic-roughness maps as data = plain box chain, no sRGB round trip; colour maps
    with setSrgbFlag(true) = dxgiFormat 99.  Defaults are restored afterwards."""
    img = synth.make("metal_rough", 128, 64, 5)
    col = synth.make("albedo", 64, 64, 1)
    base_mr = tex.encode_rgba8(img, tex.KIND_METALLIC_ROUGHNESS)
    base_col = tex.encode_rgba8(col, tex.KIND_COLOR)
    tex.shutdown()
    try:
        tex.init(linear_data_maps=True, srgb_flag=True)
        got = tex.encode_rgba8(img, tex.KIND_METALLIC_ROUGHNESS)
        want = oracle.generate_texture(img, oracle.KIND_NORMAL, oracle.FMT_BC7, oracle.Q_NORMAL, oracle.NORMALIZE_OFF)
        assert (got[148:] == want[148:]).all()
        lv = tex.mips_rgba8(img, tex.KIND_METALLIC_ROUGHNESS)
        ref = oracle.mip_chain(img, oracle.KIND_NORMAL, oracle.NORMALIZE_OFF)
        assert all((a == b).all() for a, b in zip(lv, ref))  # exact integer box sums
        assert int.from_bytes(got[128:132].tobytes(), "little") == 98  # data maps stay UNORM
        c = tex.encode_rgba8(col, tex.KIND_COLOR)
        assert int.from_bytes(c[128:132].tobytes(), "little") == 99 and (c[148:] == base_col[148:]).all()
        if oracle.have_ref_dds():
            assert oracle.ref_dds_read(c.tobytes())["format"] == 99
    finally:
        tex.shutdown()
        tex.init()
    assert (tex.encode_rgba8(img, tex.KIND_METALLIC_ROUGHNESS) == base_mr).all()
    assert (tex.encode_rgba8(col, tex.KIND_COLOR) == base_col).all()


# the product built with -DROSY_TEX_TEST_HOOKS (`make -C rosy_b200/csrc hooks`): the only library that reads the
# fault-injection variables; the tests that need them run it in a process of their own
HOOKS_LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rosy_b200", "csrc",
                         "librosy_tex_cuda_hooks.so")


def _run_py(code, env_extra):
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, **env_extra)
    out = subprocess.run([sys.executable, "-c", f"import sys; sys.path.insert(0, {root!r})\n" + code], env=env,
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    return out.stdout


def test_unexpected_shared_window_takes_generic_path():
    """the fused mip kernel's table layout depends on where dynamic shared memory starts.  The
    library probes that per device and, when it differs (faked here), takes the per-level kernels instead of
    trapping: results stay within the 1 LSB bar and the context stays usable."""
    code = """
import numpy as np
from rosy_b200 import texture as T, synth as S
import oracle as O
img = S.albedo(512, 512, 7)
for _ in range(2):
    got = T.mips_rgba8(img, T.KIND_COLOR)
want = O.mip_chain(img, O.KIND_SRGB)
print(max(int(np.abs(g.astype(int) - w.astype(int)).max()) for g, w in zip(got, want)), T.kernel_launches())
"""
    worst, launches = map(int, _run_py(code, {"ROSY_TEX_LIB": HOOKS_LIB, "ROSY_TEX_FAKE_WINDOW_MISMATCH": "1"}).split()[-2:])
    assert worst <= 1
    assert launches >= 2 * 9, "the per-level kernels must have run (one launch per level)"
    worst, launches = map(int, _run_py(code, {}).split()[-2:])
    assert worst <= 1 and launches == 2, "the fused kernel is one launch per chain"
    # the product library has no such hook: the variable changes nothing there
    worst, launches = map(int, _run_py(code, {"ROSY_TEX_FAKE_WINDOW_MISMATCH": "1"}).split()[-2:])
    assert worst <= 1 and launches == 2
    # a fused launch that refuses to run at run time (faked: the flag it would set is set after the first
    # fused launch) costs that texture one re-run on the per-level kernels and moves the device to them for good -
    # no error, nothing sticky: 1 refused fused launch + 9 per-level launches, then 9 for the second chain
    worst, launches = map(int, _run_py(code, {"ROSY_TEX_LIB": HOOKS_LIB, "ROSY_TEX_FAKE_LAYOUT_FAULT": "1"}).split()[-2:])
    assert worst <= 1 and launches == 1 + 9 + 9


def test_device_fault_requeues_textures_on_the_remaining_workers():
    """a worker that sees a device-level failure retires and its textures are finished by the others;
    bytes are those of the fault-free run.  A fault on the only worker left is reported (reference: any failure
    aborts the asset, Gltf.cpp:225-229)."""
    code = """
import numpy as np, hashlib
from rosy_b200 import texture as T, synth as S
imgs = [(S.albedo(256, 256, i), T.KIND_COLOR) for i in range(5)] + [(S.normal(1024, 1024, 9), T.KIND_NORMAL_MAP)]
q = T.Pending()
for img, kind in imgs:
    q.submit(img, kind)
outs = q.wait_all()
live, requeued = T.queue_status()
print(live, requeued, hashlib.sha256(b"".join(o.tobytes() for o in outs)).hexdigest())
"""
    live, requeued, clean = _run_py(code, {}).split()[-3:]
    assert (int(live), int(requeued)) == (2, 0)  # one B200, two workers, nothing went wrong
    live, requeued, faulty = _run_py(code, {"ROSY_TEX_LIB": HOOKS_LIB, "ROSY_TEX_FAULT_SLOT": "0"}).split()[-3:]
    assert clean == faulty
    # the degraded pool is visible: wait_all returned 0, but one worker is gone and its textures were handed over
    assert int(live) == 1 and int(requeued) >= 1
