"""Runs the fused mip (+ BC5 / BC4) kernels of a bounds-checked build (`make -C rosy_b200/csrc checked`) over single and
batched launches, every image kind, fused-encode and plain, small and 4096^2 / non-square shapes, and prints how many
stores fell outside their buffers (must be 0).  With ROSY_TEX_CHECK_SELFTEST=1 the build pretends the chain buffer is
4 KB shorter: the count must then be positive, which shows that the checks are live.  Stand-in for compute-sanitizer
memcheck, which is closed on the measurement pool (profiles/r2d_compute_sanitizer_closed.txt).

    ROSY_TEX_LIB=rosy_b200/csrc/librosy_tex_cuda_checked.so python tools/checked_run.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rosy_b200 import synth, texture as T  # noqa: E402


def main():
    T.lib()
    assert T.debug_faults() >= 0, "not a checked build: set ROSY_TEX_LIB to librosy_tex_cuda_checked.so"
    n = 0
    for (w, h) in [(64, 64), (192, 320), (1024, 512), (2048, 2048), (4096, 64), (64, 4096), (4096, 4096)]:
        for name, idx, kind, fmt, normalize in [("albedo", 0, T.KIND_COLOR, T.FORMAT_BC7, -1), ("albedo", 7, T.KIND_COLOR, T.FORMAT_BC5, -1),
                                                ("normal", 1, T.KIND_NORMAL_MAP, T.FORMAT_BC5, 1), ("normal", 4, T.KIND_NORMAL_MAP, T.FORMAT_BC4, 0),
                                                ("metal_rough", 2, T.KIND_METALLIC_ROUGHNESS, T.FORMAT_BC4, -1), ("normal", 3, T.KIND_NORMAL_MAP, T.FORMAT_BC7, 1)]:
            if w * h > 2048 * 2048 and fmt == T.FORMAT_BC7:
                q = 0
            else:
                q = 1
            img = synth.make(name, w, h, idx)
            T.encode_rgba8(img, kind, fmt, q, normalize=normalize)
            T.mips_rgba8(img, kind, normalize=normalize)
            n += 2
    pend = T.Pending()  # the work queue with a batch of small textures of mixed kinds and formats
    for i in range(24):
        kind, fmt = [(T.KIND_COLOR, T.FORMAT_BC7), (T.KIND_NORMAL_MAP, T.FORMAT_BC5), (T.KIND_METALLIC_ROUGHNESS, T.FORMAT_BC4)][i % 3]
        name = ["albedo", "normal", "metal_rough"][i % 3]
        pend.submit(synth.make(name, 64 << (i % 3), 64 << ((i + 1) % 3), i), kind, fmt, 0)
    pend.wait_all()
    faults = T.debug_faults()
    print(f"checked_run: {n} synchronous calls + a queued batch of 24, kernels launched {T.kernel_launches()}, out-of-bounds stores: {faults}"
          f"{' (self-test: the chain buffer is declared 4 KB short)' if os.environ.get('ROSY_TEX_CHECK_SELFTEST') else ''}")
    T.shutdown()
    return faults


if __name__ == "__main__":
    f = main()
    want_positive = bool(os.environ.get("ROSY_TEX_CHECK_SELFTEST"))
    sys.exit(0 if (f > 0) == want_positive else 1)
