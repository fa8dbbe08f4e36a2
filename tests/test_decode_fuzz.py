"""The image decoders (PNG incl. the own inflate and the SSE2 unfilter, JPEG, TGA) on seeded mutations of the golden
files, under AddressSanitizer + UBSan: corrupt files may decode or be rejected, never touch foreign memory."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def test_decoders_survive_mutated_files(tmp_path):
    exe = str(tmp_path / "decode_fuzz")
    csrc = os.path.join(ROOT, "rosy_b200", "csrc")
    base = ["g++", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-std=c++17",
            "-I" + os.path.join(ROOT, "include")]
    # each decoder source with the flags rosy_b200/csrc/Makefile gives it: -fwrapv for jpeg_decode.cpp ONLY (its inverse
    # DCT wraps by design), so that UBSan still sees signed overflow in the PNG / TGA / PNM size arithmetic and the inflate
    objs = []
    for f, extra in (("image_io.cpp", []), ("inflate_fast.cpp", []), ("jpeg_decode.cpp", ["-fwrapv"]),
                     (os.path.join(ROOT, "tools", "decode_fuzz.cpp"), [])):
        obj = str(tmp_path / (os.path.basename(f) + ".o"))
        r = subprocess.run(base + extra + ["-c", os.path.join(csrc, f), "-o", obj], capture_output=True, text=True)
        if r.returncode != 0 and "sanitize" in r.stderr:
            pytest.skip("no sanitizer runtime on this host")
        assert r.returncode == 0, r.stderr[-3000:]
        objs.append(obj)
    r = subprocess.run(["g++", "-fsanitize=address,undefined", "-o", exe] + objs + ["-lz"], capture_output=True, text=True)
    if r.returncode != 0 and "sanitize" in r.stderr:
        pytest.skip("no sanitizer runtime on this host")
    assert r.returncode == 0, r.stderr[-3000:]
    seeds = []
    for npz, prefix, ext in (("stb_image_png.npz", "png_", ".png"), ("stb_image_jpeg.npz", "jpg_", ".jpg"), ("stb_image_tga.npz", "tga_", ".tga")):
        g = np.load(os.path.join(GOLD, npz))
        for k in g.files:
            if k.startswith(prefix):
                p = tmp_path / (k + ext)
                p.write_bytes(g[k].tobytes())
                seeds.append(str(p))
    assert len(seeds) > 50
    env = dict(os.environ, ASAN_OPTIONS="allocator_may_return_null=1:max_allocation_size_mb=4096")
    r = subprocess.run([exe, "120"] + seeds, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-4000:])
    assert "decode_fuzz:" in r.stdout
    print(r.stdout.strip())
