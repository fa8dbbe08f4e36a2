"""Mip kernel timing by launch size (development aid): one texture per launch at 1024 / 2048 / 4096, and batches of
8 and 64 x 2048^2, colour and normal maps; CUDA events around back-to-back launches (launch latency overlaps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import numpy as np, torch
from rosy_b200 import texture as T, synth

HBM = 6452.5


def one(kind_name, kind, n, reps=50):
    img = synth.make(kind_name, n, n, 1)
    d = T.desc(n, n, kind, 0, 1, -1)
    dev = torch.device("cuda:0")
    rgba = torch.from_numpy(img).to(dev)
    chain = torch.empty(T.chain_size(n, n), dtype=torch.uint8, device=dev)
    scratch = torch.zeros(T.device_scratch_size(d), dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    fn = lambda: T.mips_device(d, rgba.data_ptr(), chain.data_ptr(), scratch.data_ptr(), st)
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    alg = n * n * (5.3333 if kind != T.KIND_NORMAL_MAP else 9.3333)
    return {"kind": kind_name, "size": n, "textures": 1, "us_per_launch": round(us, 2), "GBs": round(alg / us / 1e3, 1), "frac_hbm": round(alg / us / 1e3 / HBM, 3)}


def batch(kind_name, kind, n, count, reps=20):
    dev = torch.device("cuda:0")
    d = T.desc(n, n, kind, 0, 1, -1)
    keep, entries = [], []
    base = torch.from_numpy(synth.make(kind_name, n, n, 1)).to(dev)
    for i in range(count):
        t_in = torch.roll(base, (37 * i, 11 * i), (0, 1)).contiguous()
        t_chain = torch.empty((T.chain_size(n, n) + 255) // 256 * 256, dtype=torch.uint8, device=dev)
        t_scr = torch.zeros(T.device_scratch_size(d), dtype=torch.uint8, device=dev)
        t_blk = torch.empty(256, dtype=torch.uint8, device=dev)
        keep.append((t_in, t_chain, t_scr, t_blk))
        entries.append((d, t_in.data_ptr(), t_chain.data_ptr(), t_scr.data_ptr(), t_blk.data_ptr()))
    b = T.make_batch(entries)
    t_batch = torch.zeros(T.batch_scratch_size(count), dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    fn = lambda: T.encode_batch_device(b, t_batch.data_ptr(), st, True, False)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    alg = count * n * n * (5.3333 if kind != T.KIND_NORMAL_MAP else 9.3333)
    return {"kind": kind_name, "size": n, "textures": count, "us_per_launch": round(us, 2), "GBs": round(alg / us / 1e3, 1), "frac_hbm": round(alg / us / 1e3 / HBM, 3)}


if __name__ == "__main__":
    T.lib()
    for kn, k in (("albedo", T.KIND_COLOR), ("normal", T.KIND_NORMAL_MAP)):
        for n in (1024, 2048, 4096):
            print(json.dumps(one(kn, k, n)), flush=True)
        for c in (8, 64):
            print(json.dumps(batch(kn, k, 2048, c)), flush=True)
