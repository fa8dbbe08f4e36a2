"""Small pass over every kernel for compute-sanitizer (memcheck / racecheck / synccheck): the fused TMA mip kernel
(colour, cut-out alpha, normal; single and batched), the generic per-level mip kernels (ragged extent), the BC7
kernel at its three CTA sizes and three tiers, BC5 / BC4, and the work queue with a batch of small textures.
Outputs are compared with the CPU oracle so that a sanitizer-clean run is also a correct one.

    compute-sanitizer --tool racecheck python tools/sanitize_smoke.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as O  # noqa: E402
from rosy_b200 import synth, texture as T  # noqa: E402


def main():
    T.lib()
    cases = [("albedo", 0, 256, 256, T.KIND_COLOR, T.FORMAT_BC7, 1), ("albedo", 7, 128, 128, T.KIND_COLOR, T.FORMAT_BC7, 2),
             ("normal", 1, 256, 128, T.KIND_NORMAL_MAP, T.FORMAT_BC5, 1), ("metal_rough", 2, 128, 128, T.KIND_METALLIC_ROUGHNESS, T.FORMAT_BC4, 1),
             ("albedo", 3, 100, 60, T.KIND_COLOR, T.FORMAT_BC7, 0), ("normal", 4, 128, 128, T.KIND_NORMAL_MAP, T.FORMAT_BC7, 0)]
    for name, idx, w, h, kind, fmt, q in cases:
        img = synth.make(name, w, h, idx)
        okind = O.KIND_NORMAL if kind == T.KIND_NORMAL_MAP else O.KIND_SRGB
        got = T.mips_rgba8(img, kind)
        want = O.mip_chain(img, okind)
        worst = max(int(np.abs(g.astype(int) - w_.astype(int)).max()) for g, w_ in zip(got, want))
        assert worst <= 1, (name, worst)
        dds = T.encode_rgba8(img, kind, fmt, q)
        off = 148
        for lvl in got:  # the encoder on the GPU's own levels must equal the oracle's encoder on the same pixels
            lh, lw = lvl.shape[:2]
            n = ((lw + 3) // 4) * ((lh + 3) // 4) * (8 if fmt == T.FORMAT_BC4 else 16)
            assert (dds[off : off + n] == O.encode_image(lvl, fmt, q)).all(), (name, lw, lh)
            off += n
    big = synth.albedo(1024, 1024, 2)  # 87 383 blocks: the 640-thread kernel; 256^2 above: the 256-thread kernel
    T.encode_rgba8(big, T.KIND_COLOR, T.FORMAT_BC7, 0)
    big2 = synth.albedo(2048, 1024, 5)  # 174 763 blocks: the 512-thread kernel
    T.encode_rgba8(big2, T.KIND_COLOR, T.FORMAT_BC7, 0)
    q = T.Pending()
    small = [(synth.albedo(64, 64, i), T.KIND_COLOR) for i in range(6)] + [(synth.normal(128, 64, 9), T.KIND_NORMAL_MAP)]
    for img, kind in small:
        q.submit(img, kind)
    outs = q.wait_all()
    for (img, kind), out in zip(small, outs):
        assert (out == T.encode_rgba8(img, kind)).all()
    print("sanitize_smoke ok, kernels launched:", T.kernel_launches())
    T.shutdown()


if __name__ == "__main__":
    main()
