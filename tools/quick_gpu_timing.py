"""Scratch timing of the device-resident path (development aid; bench.py is the contract)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rosy_b200 import texture as T, synth

def run(kind_name, kind, fmt, q, n=4096, reps=5):
    img = synth.make(kind_name, n, n, 1)
    d = T.desc(n, n, kind, fmt, q, -1)
    dev = torch.device("cuda:0")
    rgba = torch.from_numpy(img).to(dev)
    chain = torch.empty(T.chain_size(n, n), dtype=torch.uint8, device=dev)
    scratch = torch.zeros(T.device_scratch_size(d), dtype=torch.uint8, device=dev)
    blocks = torch.empty(T.dds_size(d) - 148, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    for name, fn in (("mips", lambda: T.mips_device(d, rgba.data_ptr(), chain.data_ptr(), scratch.data_ptr(), st)),
                     ("blocks", lambda: T.blocks_device(d, rgba.data_ptr(), chain.data_ptr(), blocks.data_ptr(), st))):
        for _ in range(2): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(f"{kind_name:8s} fmt{fmt} q{q} {n}x{n} {name:6s} {ms:9.3f} ms  {n*n/ms/1e3:10.1f} Mpix/s", flush=True)

if __name__ == "__main__":
    for mode in ("fused", "generic"):
        os.environ["ROSY_TEX_MIPS"] = mode
    run("albedo", 0, 0, 0); run("albedo", 0, 0, 1); run("albedo", 0, 0, 2, reps=2)
    run("normal", 1, 0, 1); run("normal", 1, 1, 1); run("metal_rough", 2, 2, 1)
