#!/usr/bin/env python
"""File-to-file throughput of the work queue (decode pool + pinned staging + device
slots + writer): N image files (PNG and JPEG) -> .dds next to them, through rosy_tex_submit /
rosy_tex_wait_all.  Run once per I/O thread count in a fresh process:

    ROSY_TEX_IO_THREADS=1 python tools/bench_files.py ; python tools/bench_files.py
"""
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
from PIL import Image

import bench
from rosy_b200 import texture as T


def main():
    n, size = int(os.environ.get("N", "32")), int(os.environ.get("SIZE", "2048"))
    d = tempfile.mkdtemp(prefix="rosy_files_")
    try:
        imgs = bench.make_textures("albedo", size, n, 0)
        paths = []
        for i, img in enumerate(imgs):
            if i % 2:
                p = os.path.join(d, f"t{i}.jpg")
                Image.fromarray(img[..., :3]).save(p, quality=90)
            else:
                p = os.path.join(d, f"t{i}.png")
                Image.fromarray(img).save(p, compress_level=1)
            paths.append(p)
        in_bytes = sum(os.path.getsize(p) for p in paths)
        T.init(devices=[0], quality=1)
        T.submit_file(paths[0], T.KIND_COLOR)  # warm-up: tables, buffers, threads
        T.wait_all()
        t0 = time.perf_counter()
        for p in paths:
            T.submit_file(p, T.KIND_COLOR)
        T.wait_all()
        sec = time.perf_counter() - t0
        ok = all(os.path.getsize(os.path.splitext(p)[0] + ".dds") == T.dds_size(T.desc(size, size)) for p in paths)
        print(json.dumps({"files": n, "size": size, "io_threads": os.environ.get("ROSY_TEX_IO_THREADS", f"default ({os.cpu_count()} cores)"),
                          "seconds": sec, "mpix_s": n * size * size / 1e6 / sec, "input_MB": in_bytes / 1e6, "all_written": ok}))
        T.shutdown()
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
