#!/usr/bin/env python
"""In-process multi-GPU check of the work queue (the Packager case: ONE process, every visible B200):
the same textures through rosy_tex_submit_rgba8 / wait_all on 1 device and on all devices; outputs must
be byte-identical, throughput should scale.  Run under `gpurun --gpus N`."""
import ctypes as C
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

import bench
from rosy_b200 import texture as T


def run(devices, texs, d, steps=3):
    T.init(devices=devices, quality=1)
    L = T.lib()
    n, dds = len(texs), T.dds_size(d)
    hin = [L.rosy_tex_alloc_host(t.nbytes) for t in texs]
    hout = [L.rosy_tex_alloc_host(dds) for _ in texs]
    lens = [C.c_size_t() for _ in texs]
    for p, t in zip(hin, texs):
        C.memmove(p, t.ctypes.data, t.nbytes)

    def step():
        for i in range(n):
            assert L.rosy_tex_submit_rgba8(C.byref(d), hin[i], hout[i], dds, C.byref(lens[i])) == 0
        assert L.rosy_tex_wait_all() == 0, L.rosy_tex_last_error()

    step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    sec = (time.perf_counter() - t0) / steps
    h = hashlib.sha256()
    for p in hout:
        h.update(C.string_at(p, dds))
    for p in hin + hout:
        L.rosy_tex_free_host(p)
    ndev = T.device_count()
    T.shutdown()
    return sec, h.hexdigest(), ndev


def main():
    size, n = int(os.environ.get("SIZE", "2048")), int(os.environ.get("N", "128"))
    texs = bench.make_textures("albedo", size, n, 0)
    d = T.desc(size, size, T.KIND_COLOR, T.FORMAT_BC7, 1, -1)
    ngpu = torch.cuda.device_count()
    out = {"textures": n, "size": size, "visible_gpus": ngpu}
    s1, h1, _ = run([0], texs, d)
    out["one_device_mpix_s"] = n * size * size / 1e6 / s1
    sa, ha, nd = run(list(range(ngpu)), texs, d)
    out["all_devices"] = nd
    out["all_devices_mpix_s"] = n * size * size / 1e6 / sa
    out["identical_bytes"] = h1 == ha
    out["speedup"] = s1 / sa
    print(json.dumps(out))


if __name__ == "__main__":
    main()
