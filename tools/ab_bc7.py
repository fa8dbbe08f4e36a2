"""A/B timing of BC7 kernel builds: ROSY_TEX_LIB=<lib> python tools/ab_bc7.py  (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rosy_b200 import texture as T, synth
import oracle as O

def run(kind_name, kind, q, n=2048, reps=10):
    img = synth.make(kind_name, n, n, 1)
    d = T.desc(n, n, kind, 0, q, -1)
    dev = torch.device("cuda:0")
    rgba = torch.from_numpy(img).to(dev)
    chain = torch.empty(T.chain_size(n, n), dtype=torch.uint8, device=dev)
    scratch = torch.zeros(T.device_scratch_size(d), dtype=torch.uint8, device=dev)
    blocks = torch.empty(T.dds_size(d) - 148, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    T.mips_device(d, rgba.data_ptr(), chain.data_ptr(), scratch.data_ptr(), st)
    fn = lambda: T.blocks_device(d, rgba.data_ptr(), chain.data_ptr(), blocks.data_ptr(), st)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # parity spot check on a crop (bit exact vs the oracle)
    crop = np.ascontiguousarray(img[:64, :64])
    ok = (T.compress_level(crop, 0, q) == O.encode_image(crop, 0, q)).all()
    print(f"NT={os.environ.get('ROSY_TEX_BC7_NT', 'auto'):5s} {os.path.basename(T.LIB_PATH):12s} {kind_name:8s} q{q} {n}x{n} blocks {ms:8.3f} ms {n*n/ms/1e3:9.1f} Mpix/s exact={ok}", flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "sizes":  # single-texture launches: the CTA size choice (ROSY_TEX_BC7_NT forces one)
        for n in (256, 512, 1024, 2048, 4096):
            run("albedo", 0, 1, n=n, reps=20 if n <= 1024 else 5)
    elif len(sys.argv) > 1 and sys.argv[1] == "one":
        run("albedo", 0, 1)
    else:
        run("albedo", 0, 1); run("albedo", 0, 0); run("albedo", 0, 2, reps=3); run("normal", 1, 1)
