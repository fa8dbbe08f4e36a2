#!/usr/bin/env python
"""In-process multi-GPU measurement of the work queue - the product's multi-GPU path (north_star: "a per-texture
work queue with no collectives"; the Packager case: ONE process, every visible B200).

For each workload the same textures go through rosy_tex_submit_rgba8 / rosy_tex_wait_all on 1 device and on all
visible devices; outputs must be byte-identical, throughput should scale.  Workloads (BASELINE.json):
  c3        configs[3]  16 x 8192^2 (8 albedo + 8 normal), BC7 normal: 2 textures per device on 8 GPUs, so the
                        largest-first pop order matters
  c2        configs[2]  70 textures 1K-4K mixed albedo / normal / metal-rough, BC7 normal
  c4fast / c4normal / c4high   configs[4] slices: N x 2048^2 albedo at the three quality tiers
  small     4096 x 256^2 albedo (batched small textures)
Submission runs on SUBMIT_THREADS host threads (the entry point is thread-safe; default 4).

    gpurun --gpus 4 -- 'python tools/bench_queue_multi.py c3 c2 c4fast c4normal c4high > gpurun_out/queue_multi.jsonl'
"""
import ctypes as C
import hashlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

import bench
from rosy_b200 import texture as T

KIND = {"albedo": T.KIND_COLOR, "normal": T.KIND_NORMAL_MAP, "metal_rough": T.KIND_METALLIC_ROUGHNESS}


def big_from_base(base, size, variant):
    """size^2 texture out of a smaller seeded base: tiles of the base, each rolled / flipped differently, so every
    4x4 block differs while start-up stays at seconds (an 8192^2 noise stack takes a minute per texture on the host)"""
    b = base.shape[0]
    k = size // b
    out = np.empty((size, size, 4), dtype=np.uint8)
    for ty in range(k):
        for tx in range(k):
            v = (ty * k + tx) * 7 + variant * 13 + 1
            t = np.roll(base, (37 * v, 53 * v), axis=(0, 1))
            if v & 1:
                t = t[::-1]
            if v & 2:
                t = t[:, ::-1]
            out[ty * b:(ty + 1) * b, tx * b:(tx + 1) * b] = t
    return out


def workload(name):
    """-> (description, [(image, desc)])"""
    if name == "c3":
        bases = {k: bench.make_textures(k, 2048, 1, 0)[0] for k in ("albedo", "normal")}
        texs = []
        for i in range(16):
            k = "albedo" if i < 8 else "normal"
            texs.append((big_from_base(bases[k], 8192, i), T.desc(8192, 8192, KIND[k], T.FORMAT_BC7, 1, -1)))
        return "configs[3]: 16 x 8192^2 (8 albedo + 8 normal), BC7 normal", texs
    if name == "c2":
        kinds = ["albedo"] * 24 + ["normal"] * 24 + ["metal_rough"] * 22
        sizes = [1024] * 30 + [2048] * 30 + [4096] * 10
        rng = np.random.RandomState(0x524F5359 & 0x7FFFFFFF)
        rng.shuffle(sizes)
        pools, taken, texs = {}, {}, []
        for k, s in zip(kinds, sizes):
            pools.setdefault((k, s), bench.make_textures(k, s, sum(1 for kk, ss in zip(kinds, sizes) if (kk, ss) == (k, s)), 0))
            i = taken.get((k, s), 0)
            taken[(k, s)] = i + 1
            texs.append((pools[(k, s)][i], T.desc(s, s, KIND[k], T.FORMAT_BC7, 1, -1)))
        return "configs[2]: Sponza-scale set, 70 textures 1K-4K mixed, BC7 normal", texs
    if name in ("c4fast", "c4normal", "c4high"):
        q = {"c4fast": 0, "c4normal": 1, "c4high": 2}[name]
        n = int(os.environ.get("N", "128" if q < 2 else "64"))
        imgs = bench.make_textures("albedo", 2048, n, 0)
        return f"configs[4] slice: {n} x 2048^2 albedo, BC7 {['fast', 'normal', 'high'][q]}", [
            (im, T.desc(2048, 2048, T.KIND_COLOR, T.FORMAT_BC7, q, -1)) for im in imgs]
    if name == "small":
        imgs = bench.make_textures("albedo", 256, 4096, 0)
        return "4096 x 256^2 albedo, BC7 normal (batched small textures)", [
            (im, T.desc(256, 256, T.KIND_COLOR, T.FORMAT_BC7, 1, -1)) for im in imgs]
    raise SystemExit(f"unknown workload {name}")


def run(devices, texs, steps, submit_threads):
    T.init(devices=devices, quality=1)
    L = T.lib()
    n = len(texs)
    sizes = [T.dds_size(d) for _, d in texs]
    hin = [L.rosy_tex_alloc_host(im.nbytes) for im, _ in texs]
    hout = [L.rosy_tex_alloc_host(s) for s in sizes]
    assert all(hin) and all(hout), "pinned allocation failed"
    lens = [C.c_size_t() for _ in texs]
    for p, (im, _) in zip(hin, texs):
        C.memmove(p, im.ctypes.data, im.nbytes)

    def submit_range(lo, hi):
        for i in range(lo, hi):
            assert L.rosy_tex_submit_rgba8(C.byref(texs[i][1]), hin[i], hout[i], sizes[i], C.byref(lens[i])) == 0

    def step():
        k = max(1, min(submit_threads, n))
        if k == 1:
            submit_range(0, n)
        else:  # the entry point is thread-safe: several importer threads may feed the queue
            th = [threading.Thread(target=submit_range, args=(n * j // k, n * (j + 1) // k)) for j in range(k)]
            for t in th:
                t.start()
            for t in th:
                t.join()
        assert L.rosy_tex_wait_all() == 0, L.rosy_tex_last_error()

    step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    sec = (time.perf_counter() - t0) / steps
    h = hashlib.sha256()
    for p, s in zip(hout, sizes):
        h.update(C.string_at(p, s))
    for p in hin + hout:
        L.rosy_tex_free_host(p)
    ndev = T.device_count()
    T.shutdown()
    return sec, h.hexdigest(), ndev


def main():
    names = sys.argv[1:] or ["c4normal"]
    ngpu = torch.cuda.device_count()
    st = int(os.environ.get("SUBMIT_THREADS", "4"))
    steps = int(os.environ.get("STEPS", "3"))
    for name in names:
        desc, texs = workload(name)
        mpix = sum(im.shape[0] * im.shape[1] for im, _ in texs) / 1e6
        out = {"workload": desc, "textures": len(texs), "mpix_level0": mpix, "visible_gpus": ngpu, "submit_threads": st}
        s1, h1, _ = run([0], texs, steps, st)
        out["one_device_mpix_s"] = mpix / s1
        if ngpu > 1:
            sa, ha, nd = run(list(range(ngpu)), texs, steps, st)
            out["all_devices"] = nd
            out["all_devices_mpix_s"] = mpix / sa
            out["identical_bytes"] = h1 == ha
            out["speedup"] = s1 / sa
            out["efficiency"] = s1 / sa / nd
        print(json.dumps(out), flush=True)
        del texs


if __name__ == "__main__":
    main()
