# one-off: the largest extent the ABI accepts (32768^2 = 4 GiB of RGBA8): addressing beyond 32 bits
import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
from rosy_b200 import texture as T, synth as S
T.init(devices=[0], quality=0)
base = S.normal(512, 512, 1)
small = T.encode_rgba8(base, T.KIND_NORMAL_MAP, T.FORMAT_BC5, 1)
for (w, h) in [(32768, 16384), (32768, 32768)]:
    big = np.ascontiguousarray(np.tile(base, (h // 512, w // 512, 1)))
    t0 = time.time()
    got = T.encode_rgba8(big, T.KIND_NORMAL_MAP, T.FORMAT_BC5, 1)
    dt = time.time() - t0
    off_s, off_b, ok = 148, 148, True
    for k in range(8):
        sb = (512 >> k) // 4
        bx, by = (w >> k) // 4, (h >> k) // 4
        a = small[off_s: off_s + sb * sb * 16].reshape(sb, sb, 16)
        b = got[off_b: off_b + bx * by * 16].reshape(by, bx, 16)
        same = bool((b == np.tile(a, (h // 512, w // 512, 1))).all())
        ok &= same
        off_s += sb * sb * 16; off_b += bx * by * 16
    print(w, h, "periodic levels 0..7:", ok, "seconds", round(dt, 2), flush=True)
