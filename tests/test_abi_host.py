"""CPU suite, part 2: the C-ABI library (the product) loads here, exports every symbol that
include/rosy_tex_cuda.h declares, its host-only entry points work, and every compute entry point
fails LOUDLY (ROSY_TEX_ERR_NO_DEVICE) when there is no sm_100 device - there is no CPU fallback."""
import ctypes as C
import io
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _declared_symbols():
    with open(os.path.join(ROOT, "include", "rosy_tex_cuda.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"ROSY_TEX_API\s+[\w\s\*]+?\b(rosy_tex_\w+)\s*\(", text)))


def _no_gpu():
    import torch

    return not torch.cuda.is_available()


def test_library_exports_every_declared_symbol(tex):
    L = tex.lib()
    names = _declared_symbols()
    assert len(names) >= 25, names
    for n in names:
        assert hasattr(L, n), f"{n} is declared in include/rosy_tex_cuda.h but not exported"


def test_no_oracle_in_the_product_path():
    """the product never imports, links or calls the oracle"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "rosy_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cpp", ".h", ".cuh", "Makefile")):
                with open(os.path.join(dirpath, fn), errors="replace") as f:
                    text = f.read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), fn
                assert "rosy_oracle.h" not in text and "liboracle" not in text, fn
    import subprocess

    out = subprocess.run(["ldd", os.path.join(ROOT, "rosy_b200", "csrc", "librosy_tex_cuda.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out


def test_geometry_helpers_match_oracle(tex, oracle):
    for (w, h) in [(1024, 1024), (4096, 4096), (8192, 8192), (1, 1), (5, 3), (256, 64), (100, 60), (2048, 2048)]:
        assert tex.count_mipmaps(w, h) == oracle.count_mipmaps(w, h)
        assert tex.chain_size(w, h) == 4 * oracle.chain_pixels(w, h)
        for l in range(tex.count_mipmaps(w, h)):
            assert tex.level_extent(w, h, l) == oracle.level_extent(w, h, l)
        for fmt in (0, 1, 2):
            d = tex.desc(w, h, tex.KIND_COLOR, fmt, 1, 1)
            assert tex.dds_size(d) == oracle.dds_size(w, h, fmt)
            assert (tex.dds_header(d) == oracle.dds_header(w, h, oracle.count_mipmaps(w, h), fmt)).all()
    assert tex.count_mipmaps(0, 4) == 0 and tex.chain_size(0, 0) == 0


def test_dds_header_matches_reference_reader_vectors(tex):
    with open(os.path.join(GOLD, "dds_reader.json")) as f:
        for c in json.load(f):
            d = tex.desc(c["w"], c["h"], tex.KIND_NORMAL_MAP, c["fmt"], 1, 1)
            assert bytes(tex.dds_header(d)).hex() == c["header_hex"]


def test_image_loader_matches_reference_stb_image_vectors(tex, tmp_path):
    g = np.load(os.path.join(GOLD, "stb_image_png.npz"))
    keys = [k[4:] for k in g.files if k.startswith("png_")]
    assert len(keys) >= 8
    for k in keys:
        p = tmp_path / (k + ".png")
        p.write_bytes(g["png_" + k].tobytes())
        got = tex.load_image(str(p))
        assert got.shape == g["rgba_" + k].shape, k
        assert (got == g["rgba_" + k]).all(), k


def test_jpeg_loader_matches_reference_stb_image_vectors(tex, tmp_path):
    """baseline / progressive / restart-interval JPEGs at 4:4:4, 4:2:2, 4:2:0 and greyscale decode to the
    bytes the reference tree's stb_image.h produces (Surface::load stand-in)"""
    g = np.load(os.path.join(GOLD, "stb_image_jpeg.npz"))
    keys = [k[4:] for k in g.files if k.startswith("jpg_")]
    assert len(keys) >= 60
    for k in keys:
        p = tmp_path / (k + ".jpg")
        p.write_bytes(g["jpg_" + k].tobytes())
        got = tex.load_image(str(p))
        assert got.shape == g["rgba_" + k].shape, k
        assert (got == g["rgba_" + k]).all(), k


def test_tga_loader_matches_reference_stb_image_vectors(tex, tmp_path):
    """every TGA flavour the reference tree's stb_image.h reads (colour-mapped 1/9 with 15/16/24/32-bit
    palettes and 8/16-bit indices, true colour 15/16/24/32, grey 8/16, raw + RLE, both origins) decodes to
    stb's bytes; the FBX route (Fbx.cpp:421-474) feeds TGA exclusively"""
    g = np.load(os.path.join(GOLD, "stb_image_tga.npz"))
    keys = [k[4:] for k in g.files if k.startswith("tga_")]
    assert len(keys) >= 24
    for k in keys:
        p = tmp_path / (k + ".tga")
        p.write_bytes(g["tga_" + k].tobytes())
        got = tex.load_image(str(p))
        assert got.shape == g["rgba_" + k].shape, k
        assert (got == g["rgba_" + k]).all(), k


def test_hostile_headers_are_rejected_before_allocation(tex, tmp_path):
    """a TGA / PNM header may not make the loader allocate (or index) past the file's payload"""
    from rosy_b200 import RosyTexError

    cases = {
        "huge_raw.tga": bytes([0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0xFF, 0xFF, 0xFF, 0xFF, 32, 0]) + b"\0" * 64,
        "huge_rle.tga": bytes([0, 0, 10, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0xFF, 0x7F, 0xFF, 0x7F, 32, 0]) + b"\xff" * 64,
        "idlen_past_end.tga": bytes([200, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 4, 0, 4, 0, 24, 0]) + b"\0" * 10,
        "no_palette.tga": bytes([0, 1, 1, 0, 0, 0, 0, 24, 0, 0, 0, 0, 4, 0, 4, 0, 8, 0]) + b"\0" * 16,
        "short_palette.tga": bytes([0, 1, 1, 0, 0, 0xFF, 0xFF, 24, 0, 0, 0, 0, 4, 0, 4, 0, 8, 0]) + b"\0" * 16,
        "huge.ppm": b"P6\n4000000000 4000000000\n255\n" + b"\0" * 32,
        "wrap.ppm": b"P6\n65536 65536\n255\n" + b"\0" * 32,
    }
    for name, data in cases.items():
        p = tmp_path / name
        p.write_bytes(data)
        with pytest.raises(RosyTexError) as e:
            tex.load_image(str(p))
        assert e.value.status == 2, name


def test_jpeg_malformed_huffman_tables_are_rejected(tex, tmp_path):
    """an over-subscribed DHT (e.g. 255 codes of length 1) used to write far past the
    9-bit look-ahead table; code lengths are now validated before any table entry is written"""
    from rosy_b200 import RosyTexError

    g = np.load(os.path.join(GOLD, "stb_image_jpeg.npz"))
    good = g["jpg_64x48_420"].tobytes()
    soi_rest = good[2:]

    def dht(counts, nsym):
        body = bytes([0x00]) + bytes(counts) + bytes(range(nsym % 256)) * (nsym // 256 + 1)
        body = body[: 17 + nsym]
        return b"\xff\xc4" + (len(body) + 2).to_bytes(2, "big") + body

    bad = {
        "len1_x255": dht([255] + [0] * 15, 255),
        "len1_x3": dht([3] + [0] * 15, 3),
        "len2_x5": dht([0, 5] + [0] * 14, 5),
        "oversub_late": dht([1, 1, 1, 1, 1, 1, 1, 1, 3] + [0] * 7, 11),
        "n256_len8": dht([0] * 7 + [255, 1] + [0] * 7, 256),
    }
    for name, seg in bad.items():
        p = tmp_path / (name + ".jpg")
        p.write_bytes(b"\xff\xd8" + seg + soi_rest)
        if name == "n256_len8":  # 255 codes of length 8 + 1 of length 9 is a legal (complete-minus-one) table
            try:
                tex.load_image(str(p))
            except RosyTexError as e:
                assert e.status == 2
            continue
        with pytest.raises(RosyTexError) as e:
            tex.load_image(str(p))
        assert e.value.status == 2, name


def test_jpeg_loader_rejects_bad_files(tex, tmp_path):
    from rosy_b200 import RosyTexError

    g = np.load(os.path.join(GOLD, "stb_image_jpeg.npz"))
    good = g["jpg_64x48_420"].tobytes()
    cases = {"truncated_header": good[:20], "no_scan": good[: good.index(b"\xff\xda")] + b"\xff\xd9",
             "arithmetic": good.replace(b"\xff\xc0", b"\xff\xc9", 1), "twelve_bit": None}
    sof = good.index(b"\xff\xc0")
    cases["twelve_bit"] = good[: sof + 4] + bytes([12]) + good[sof + 5 :]
    for name, data in cases.items():
        p = tmp_path / (name + ".jpg")
        p.write_bytes(data)
        with pytest.raises(RosyTexError) as e:
            tex.load_image(str(p))
        assert e.value.status == 2, name
    # a file cut in the middle of the entropy-coded data still decodes (missing bits read as zero), like stb
    p = tmp_path / "cut.jpg"
    p.write_bytes(good[: len(good) * 2 // 3])
    assert tex.load_image(str(p)).shape == (48, 64, 4)


def test_image_loader_formats_and_errors(tex, synth, tmp_path):
    from PIL import Image
    from rosy_b200 import RosyTexError

    img = synth.albedo(33, 17, 7)
    # interlaced + 16-bit PNGs, TGA (raw + RLE, top/bottom origin), PPM, raw .rgba8 side format
    Image.fromarray(img).save(tmp_path / "a.png")
    assert (tex.load_image(str(tmp_path / "a.png")) == img).all()
    hdr = bytes([0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 33, 0, 17, 0, 32, 0x28])
    (tmp_path / "t.tga").write_bytes(hdr + img[..., [2, 1, 0, 3]].tobytes())
    assert (tex.load_image(str(tmp_path / "t.tga")) == img).all()
    hdr_bl = bytes([0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 33, 0, 17, 0, 24, 0x00])  # bottom-left origin, BGR
    (tmp_path / "b.tga").write_bytes(hdr_bl + img[::-1, :, [2, 1, 0]].tobytes())
    got = tex.load_image(str(tmp_path / "b.tga"))
    assert (got[..., :3] == img[..., :3]).all() and (got[..., 3] == 255).all()
    Image.fromarray(img[..., :3]).save(tmp_path / "p.ppm")
    got = tex.load_image(str(tmp_path / "p.ppm"))
    assert (got[..., :3] == img[..., :3]).all() and (got[..., 3] == 255).all()
    with pytest.raises(RosyTexError) as e:
        tex.load_image(str(tmp_path / "missing.png"))
    assert e.value.status == 1
    (tmp_path / "junk.png").write_bytes(b"\x89PNG\r\n\x1a\n" + b"garbage" * 10)
    with pytest.raises(RosyTexError) as e:
        tex.load_image(str(tmp_path / "junk.png"))
    assert e.value.status == 2


def test_argument_errors(tex):
    from rosy_b200 import RosyTexError

    L = tex.lib()
    assert L.rosy_tex_generate_srgb_texture(None) == 3
    assert L.rosy_tex_generate_texture_file(b"/tmp/x.png", 9, None) == 3
    assert tex.dds_size(tex.desc(0, 16)) == 0
    with pytest.raises(RosyTexError):
        tex.dds_header(tex.desc(16, 16, 0, 7))
    assert L.rosy_tex_status_string(5).decode().startswith("no sm_100")


def test_compute_entry_points_fail_loudly_without_a_device(tex, synth, tmp_path):
    if not _no_gpu():
        pytest.skip("a GPU is present: covered by the gpu suite")
    from PIL import Image
    from rosy_b200 import RosyTexError

    img = synth.albedo(16, 16, 0)
    for fn in (lambda: tex.encode_rgba8(img), lambda: tex.mips_rgba8(img), lambda: tex.compress_level(img)):
        with pytest.raises(RosyTexError) as e:
            fn()
        assert e.value.status == 5 and "no CPU fallback" in str(e.value)
    p = tmp_path / "a.png"
    Image.fromarray(img).save(p)
    with pytest.raises(RosyTexError) as e:
        tex.generate_srgb_texture(str(p))
    assert e.value.status == 5
    assert not (tmp_path / "a.dds").exists(), "a failed call must leave no .dds behind"
    assert tex.device_count() == 0


def test_png_suite_of_the_reference_tree(tex, oracle):
    """every PNG of src/libs/stb/tests/pngsuite that the reference's stb_image.h decodes gives the same
    bytes here (only where the reference tree is present: never on the GPU box)"""
    import glob

    root = "/root/reference/src/libs/stb/tests/pngsuite"
    files = sorted(glob.glob(root + "/**/*.png", recursive=True))
    if not files or not oracle.have_ref_stb():
        pytest.skip("reference tree not present")
    from rosy_b200 import RosyTexError

    checked = 0
    for f in files:
        with open(f, "rb") as fh:
            data = fh.read()
        try:
            ref = oracle.ref_image_load(data)
        except Exception:
            ref = None
        if ref is None:  # corrupt on purpose (x*.png): must be rejected here too
            with pytest.raises(RosyTexError):
                tex.load_image(f)
            continue
        got = tex.load_image(f)
        assert got.shape == ref.shape and (got == ref).all(), f
        checked += 1
    assert checked >= 240


def test_product_library_reads_no_fault_injection_variables():
    """the hooks tests/ uses to fake a lost device or an unexpected shared-memory layout live in
    librosy_tex_cuda_hooks.so only; an environment variable cannot make the shipping library fail"""
    csrc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rosy_b200", "csrc")
    product = open(os.path.join(csrc, "librosy_tex_cuda.so"), "rb").read()
    hooks = open(os.path.join(csrc, "librosy_tex_cuda_hooks.so"), "rb").read()
    for name in (b"ROSY_TEX_FAULT_SLOT\0", b"ROSY_TEX_FAKE_WINDOW_MISMATCH\0", b"ROSY_TEX_FAKE_LAYOUT_FAULT\0"):
        assert name not in product, name
        assert name in hooks, name
