"""PSNR of the three product BC7 tiers against the oracle-only exhaustive ceiling (VERDICT r1 task 1b).

CPU only (oracle/bc7.c).  For every texture kind: the mip chain of a SIZE^2 texture is encoded at
fast / normal / high / exhaustive, decoded with the spec decoder and compared with the uncompressed level.
Reported per level and chain-weighted (mean squared error pooled over all texels of all levels).

    python tools/bc7_quality_gap.py [--size 256] [--out profiles/r2_bc7_quality_gap.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as O  # noqa: E402
from rosy_b200 import synth  # noqa: E402

KINDS = [("albedo", 0, O.KIND_SRGB), ("albedo_cutout", 7, O.KIND_SRGB), ("normal", 1, O.KIND_NORMAL),
         ("metal_rough", 2, O.KIND_SRGB), ("random", 3, O.KIND_SRGB)]
TIERS = [("fast", O.Q_FAST), ("normal", O.Q_NORMAL), ("high", O.Q_HIGH), ("exhaustive", O.Q_EXHAUSTIVE)]


def psnr_from_sse(sse, n):
    return 99.0 if sse == 0 else 10.0 * np.log10(255.0 * 255.0 * n / sse)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rows = []
    for name, idx, kind in KINDS:
        img = synth.make("albedo" if name.startswith("albedo") else name, a.size, a.size, idx)
        chain = O.mip_chain(img, kind)
        rec = {"kind": name, "size": a.size, "levels": len(chain), "tiers": {}}
        for tname, q in TIERS:
            t0 = time.time()
            sse_l, n_l = [], []
            for lvl in chain:
                h, w = lvl.shape[:2]
                dec = O.decode_image(O.encode_image(lvl, O.FMT_BC7, q), w, h)
                d = lvl.astype(np.int64) - dec.astype(np.int64)
                sse_l.append(int((d * d).sum()))
                n_l.append(int(d.size))
            rec["tiers"][tname] = {
                "psnr_per_level": [round(float(psnr_from_sse(s, n)), 3) for s, n in zip(sse_l, n_l)],
                "psnr_chain": round(float(psnr_from_sse(sum(sse_l), sum(n_l))), 3),
                "seconds": round(time.time() - t0, 2),
            }
        ex = rec["tiers"]["exhaustive"]
        for tname, _ in TIERS[:3]:
            t = rec["tiers"][tname]
            t["delta_chain_vs_exhaustive"] = round(t["psnr_chain"] - ex["psnr_chain"], 3)
            t["delta_per_level_vs_exhaustive"] = [round(x - y, 3) for x, y in zip(t["psnr_per_level"], ex["psnr_per_level"])]
        rows.append(rec)
        print(f"{name:14s} chain PSNR  " + "  ".join(f"{t}: {rec['tiers'][t]['psnr_chain']:.2f}" for t, _ in TIERS) +
              f"   normal - exhaustive = {rec['tiers']['normal']['delta_chain_vs_exhaustive']:+.3f} dB", flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump({"tool": "tools/bc7_quality_gap.py", "spec": "v2", "threads": O.max_threads(), "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
