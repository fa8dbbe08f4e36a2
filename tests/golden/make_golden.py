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

# ---- JPEG: decoded by the reference tree's stb_image.h (pins rosy_b200/csrc/jpeg_decode.cpp) ----
rng = np.random.RandomState(7)
jout = {}
for (w, h) in [(64, 48), (37, 53), (1, 1), (2, 3), (17, 9), (33, 16)]:
    img = synth.make("albedo", w, h, 3)[..., :3] if w >= 8 and h >= 8 else rng.randint(0, 255, (h, w, 3), dtype=np.uint8)
    variants = {
        "444": dict(subsampling=0), "422": dict(subsampling=1), "420": dict(subsampling=2),
        "420_prog": dict(subsampling=2, progressive=True), "444_prog_q95": dict(subsampling=0, progressive=True, quality=95),
        "422_prog_q50": dict(subsampling=1, progressive=True, quality=50), "420_rst": dict(subsampling=2, restart_marker_blocks=3),
        "420_prog_rst": dict(subsampling=2, progressive=True, restart_marker_rows=1), "gray": dict(gray=True),
        "gray_prog": dict(gray=True, progressive=True), "420_opt_q30": dict(subsampling=2, optimize=True, quality=30),
    }
    for name, kw in variants.items():
        kw = dict(kw)
        gray = kw.pop("gray", False)
        im = Image.fromarray(img).convert("L") if gray else Image.fromarray(img)
        buf = io.BytesIO()
        im.save(buf, "JPEG", **({"quality": 85} | kw))
        data = np.frombuffer(buf.getvalue(), dtype=np.uint8)
        dec = O.ref_image_load(data.tobytes())
        assert dec is not None, name
        jout[f"jpg_{w}x{h}_{name}"] = data
        jout[f"rgba_{w}x{h}_{name}"] = dec
np.savez_compressed(os.path.join(HERE, "stb_image_jpeg.npz"), **jout)
print("golden vectors written:", sorted(os.listdir(HERE)))
