"""rosy_b200/csrc/inflate_fast.cpp (the PNG loader's DEFLATE decoder) against zlib: seeded streams of every level,
strategy and framing must inflate to the same bytes; wrong sizes must be refused; truncated and bit-flipped streams
must not touch memory they do not own (the harness is built with AddressSanitizer + UBSan)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_inflate_fast_matches_zlib_and_survives_corruption(tmp_path):
    exe = str(tmp_path / "inflate_harness")
    cmd = ["g++", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-std=c++17", "-o", exe,
           os.path.join(ROOT, "tools", "inflate_harness.cpp"), os.path.join(ROOT, "rosy_b200", "csrc", "inflate_fast.cpp"), "-lz"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and "sanitize" in r.stderr:
        cmd = [c for c in cmd if "sanitize" not in c]
        r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([exe, "600", "notiming"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-1000:], r.stderr[-3000:])
    assert "fails 0" in r.stdout
