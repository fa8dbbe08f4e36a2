"""Generates the committed golden vectors under tests/golden/ (run in the build container only).

Sources of truth, none of them ours:
  * the reference's own vendored code compiled as is into oracle/_ref (recipe: oracle/Makefile `ref`):
      stb_dxt.h (BC4/BC5 encoder), stb_image.h (PNG decode), stb_image_resize2.h (sRGB box filter),
      dds.hpp (the Engine's DDS reader)            -- /root/reference/src/libs/...
  * Pillow's independent BCn decoder (PIL 12.2 "bcn" raw decoder) for BC7 / BC4 / BC5 blocks.
/root/reference does not exist on the GPU box, hence the vectors are committed.

    python tests/golden/make_golden.py
"""
import io
import json
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle as O  # noqa: E402
from rosy_b200 import synth  # noqa: E402

assert O.have_ref_stb() and O.have_ref_dds(), "build oracle/_ref first (make -C oracle ref)"
rng = np.random.default_rng(0x524F5359)

# ---- BC4 / BC5: the reference's stb_dxt.h ------------------------------------------------------
src = [rng.integers(0, 256, 16, dtype=np.uint8) for _ in range(400)]
src += [np.full(16, v, dtype=np.uint8) for v in (0, 1, 127, 128, 254, 255)]
src += [np.where(rng.integers(0, 2, 16) == 1, a, b).astype(np.uint8) for a, b in [(0, 255), (10, 12), (100, 107), (100, 108), (3, 200)]]
src += [np.clip(np.arange(16) * s + o, 0, 255).astype(np.uint8) for s, o in [(1, 0), (1, 120), (3, 90), (7, 0), (17, 0)]]
src += [np.clip(128 + rng.integers(-r, r + 1, 16), 0, 255).astype(np.uint8) for r in (1, 2, 3, 4, 7, 8, 9, 20) for _ in range(8)]
src = np.stack(src)
bc4 = np.stack([O.ref_bc4_block(s) for s in src])
rg = rng.integers(0, 256, (256, 32), dtype=np.uint8)
bc5 = np.stack([O.ref_bc5_block(s) for s in rg])
np.savez_compressed(os.path.join(HERE, "bc45_stb_dxt.npz"), bc4_src=src, bc4_blocks=bc4, bc5_src=rg, bc5_blocks=bc5)

# ---- BC7 / BC4 decode: Pillow ---------------------------------------------------------------------
blocks = rng.integers(0, 256, (6000, 16), dtype=np.uint8)
blocks = blocks[O.bc7_block_modes(blocks) >= 0][:4096]
per_mode = []
for m in range(8):  # make sure every mode is well represented
    b = rng.integers(0, 256, (256, 16), dtype=np.uint8)
    b[:, 0] = ((b[:, 0].astype(np.int32) & (0xFF ^ ((1 << (m + 1)) - 1))) | (1 << m)).astype(np.uint8)
    per_mode.append(b)
blocks = np.concatenate([blocks[: 4096 - 8 * 256]] + per_mode)
assert blocks.shape == (4096, 16) and (O.bc7_block_modes(blocks) >= 0).all()
pix = np.array(Image.frombytes("RGBA", (256, 256), blocks.tobytes(), "bcn", 7))
b4 = rng.integers(0, 256, (1024, 8), dtype=np.uint8)
pix4 = np.array(Image.frombytes("L", (128, 128), b4.tobytes(), "bcn", 4))
np.savez_compressed(os.path.join(HERE, "bcn_decode_pillow.npz"), bc7_blocks=blocks, bc7_rgba=pix, bc4_blocks=b4, bc4_l=pix4)

# ---- DDS container: the reference's dds::readImage ----------------------------------------------------
cases = []
for (w, h) in [(1024, 1024), (4096, 4096), (256, 64), (100, 60), (1, 1), (5, 3), (2048, 2048)]:
    for fmt in (O.FMT_BC7, O.FMT_BC5, O.FMT_BC4):
        n = O.dds_size(w, h, fmt)
        data = np.zeros(n, dtype=np.uint8)
        data[:148] = O.dds_header(w, h, O.count_mipmaps(w, h), fmt)
        info = O.ref_dds_read(data)
        cases.append({"w": w, "h": h, "fmt": fmt, "size": int(n), "header_hex": bytes(data[:148]).hex(), "parsed": info})
with open(os.path.join(HERE, "dds_reader.json"), "w") as f:
    json.dump(cases, f, indent=0)

# ---- mip level 1: stb_image_resize2 sRGB box (independent model of rows A5-A7) -------------------
img_a = synth.albedo(64, 64, 7)   # index 7: cut-out alpha
img_o = synth.albedo(96, 32, 2)
np.savez_compressed(
    os.path.join(HERE, "stbir_srgb_box.npz"),
    in_a=img_a, out_a=O.ref_resize_srgb_box(img_a), in_o=img_o, out_o=O.ref_resize_srgb_box(img_o),
)

# ---- image decode: stb_image on PNGs written by Pillow ----------------------------------------------
pngs = {}
base = synth.albedo(37, 21, 7)
variants = {
    "rgba8": Image.fromarray(base),
    "rgb8": Image.fromarray(base[..., :3]),
    "gray8": Image.fromarray(base[..., 0]),
    "graya8": Image.fromarray(base[..., [0, 3]], "LA"),
    "pal8": Image.fromarray(base[..., :3]).quantize(64),
    "rgb16": None,
}
for name, im in variants.items():
    if im is None:
        continue
    for opts in ({}, {"compress_level": 0}):
        buf = io.BytesIO()
        im.save(buf, "PNG", **opts)
        key = name + ("_stored" if opts else "")
        pngs[key] = np.frombuffer(buf.getvalue(), dtype=np.uint8)
out = {}
for k, v in pngs.items():
    dec = O.ref_image_load(v.tobytes())
    assert dec is not None, k
    out["png_" + k] = v
    out["rgba_" + k] = dec
np.savez_compressed(os.path.join(HERE, "stb_image_png.npz"), **out)
print("golden vectors written:", sorted(os.listdir(HERE)))
