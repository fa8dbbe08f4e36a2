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


# ---- BC7 ENCODER, spec v2 (frozen): the blocks oracle/bc7.c emits at the three product tiers --------
# Not an outside source of truth (NVTT is absent): this file FREEZES the search specification, so that
# oracle/bc7.c and the CUDA kernel cannot drift together unnoticed.  A deliberate spec change regenerates
# it under a new name (bc7_spec_v3.npz ...) together with tools/bc7_quality_gap.py's PSNR table.
def bc7_spec_inputs():
    rng = np.random.default_rng(0xBC7)
    srcs = [synth.albedo(96, 96, 0), synth.albedo(96, 96, 7), synth.normal(96, 96, 1), synth.metal_rough(96, 96, 2),
            synth.random_rgba(48, 48, 3), synth.gradient(96, 96)]
    srcs += O.mip_chain(synth.albedo(128, 128, 15), O.KIND_SRGB)[1:3]      # filtered levels, cut-out alpha
    srcs += O.mip_chain(synth.normal(128, 128, 4), O.KIND_NORMAL)[1:3]
    blocks = []
    for im in srcs:
        h, w = im.shape[:2]
        for by in range(h // 4):
            for bx in range(w // 4):
                blocks.append(im[4 * by : 4 * by + 4, 4 * bx : 4 * bx + 4].reshape(64))
    for _ in range(64):     # solid colours, with and without alpha
        c = rng.integers(0, 256, 4)
        c[3] = 255 if rng.integers(0, 2) else c[3]
        blocks.append(np.tile(c.astype(np.uint8), 16))
    for _ in range(128):    # two colours split by a random mask; low-variance blocks; smooth alpha ramps
        a, b = rng.integers(0, 256, 4), rng.integers(0, 256, 4)
        a[3] = b[3] = 255
        m = rng.integers(0, 2, 16).astype(bool)
        blocks.append(np.where(m[:, None], a, b).astype(np.uint8).reshape(64))
        base = rng.integers(0, 256, 4)
        blocks.append(np.clip(base + rng.integers(-3, 4, (16, 4)), 0, 255).astype(np.uint8).reshape(64))
        ramp = np.clip(base[None, :] + np.outer(np.arange(16), rng.integers(-8, 9, 4)), 0, 255).astype(np.uint8)
        blocks.append(ramp.reshape(64))
    return np.stack(blocks)


px = bc7_spec_inputs()
assert px.shape[0] >= 4096
enc = {}
for name, q in (("fast", O.Q_FAST), ("normal", O.Q_NORMAL), ("high", O.Q_HIGH)):
    enc[name] = np.stack([O.bc7_encode_block(p, q) for p in px])
assert set(O.bc7_block_modes(enc["high"].reshape(-1)).tolist()) == set(range(8)), "every BC7 mode must be covered"
np.savez_compressed(os.path.join(HERE, "bc7_spec_v2.npz"), px=px, fast=enc["fast"], normal=enc["normal"], high=enc["high"])
print("bc7_spec_v2:", px.shape[0], "blocks; modes at normal:", np.bincount(O.bc7_block_modes(enc["normal"].reshape(-1)), minlength=8).tolist(),
      "high:", np.bincount(O.bc7_block_modes(enc["high"].reshape(-1)), minlength=8).tolist())


# ---- TGA: decoded by the reference tree's stb_image.h (pins decode_tga in rosy_b200/csrc/image_io.cpp) ----
# The FBX route feeds TGA exclusively (Fbx.cpp:424).  Every type stb reads: 1/9 colour-mapped (8/16-bit
# indices; 15/16/24/32-bit palette), 2/10 true colour (15/16/24/32), 3/11 grey (8, 16 = grey + alpha),
# raw and RLE, both vertical origins, an image-id field, a palette offset.
def tga_file(type_, w, h, bpp, payload, desc=0, idlen=0, cmap=None, cmap_bits=0, cmap_first=0):
    hdr = bytearray(18)
    hdr[0] = idlen
    hdr[1] = 1 if cmap is not None else 0
    hdr[2] = type_
    if cmap is not None:
        hdr[3:5] = int(cmap_first).to_bytes(2, "little")
        hdr[5:7] = (len(cmap) // ((cmap_bits + 7) // 8)).to_bytes(2, "little")
        hdr[7] = cmap_bits
    hdr[12:14] = int(w).to_bytes(2, "little")
    hdr[14:16] = int(h).to_bytes(2, "little")
    hdr[16] = bpp
    hdr[17] = desc
    return bytes(hdr) + bytes(range(idlen)) + bytes(cmap_first) + (bytes(cmap) if cmap is not None else b"") + bytes(payload)


def tga_rle(pixels, nbytes, rng_):
    """pixels: bytes, nbytes per pixel -> RLE packets (mix of run and raw packets)"""
    px = [bytes(pixels[i : i + nbytes]) for i in range(0, len(pixels), nbytes)]
    out, i = bytearray(), 0
    while i < len(px):
        run = 1
        while i + run < len(px) and px[i + run] == px[i] and run < 128:
            run += 1
        if run >= 2:
            out += bytes([0x80 | (run - 1)]) + px[i]
            i += run
        else:
            n = min(int(rng_.integers(1, 9)), len(px) - i)
            out += bytes([n - 1]) + b"".join(px[i : i + n])
            i += n
    return bytes(out)


trng = np.random.default_rng(0x76A)
tw, th = 23, 11
base = synth.albedo(tw, th, 7)
base[:, :6] = base[:, :1]  # runs for the RLE packets
tout = {}


def tga_case(name, data):
    dec = O.ref_image_load(data)
    assert dec is not None, name
    tout["tga_" + name] = np.frombuffer(data, dtype=np.uint8)
    tout["rgba_" + name] = dec


bgr, bgra = base[..., [2, 1, 0]].tobytes(), base[..., [2, 1, 0, 3]].tobytes()
grey = base[..., 0].tobytes()
px16 = ((base[..., 0].astype(np.uint16) >> 3) << 10 | (base[..., 1].astype(np.uint16) >> 3) << 5 | (base[..., 2].astype(np.uint16) >> 3) | 0x8000).astype("<u2").tobytes()
ga = base[..., [0, 3]].tobytes()
for desc, dn in ((0x00, "bl"), (0x20, "tl"), (0x10, "br")):
    tga_case(f"t2_24_{dn}", tga_file(2, tw, th, 24, bgr, desc))
    tga_case(f"t2_32_{dn}", tga_file(2, tw, th, 32, bgra, desc | 8))
tga_case("t2_16", tga_file(2, tw, th, 16, px16, 0x20))
tga_case("t2_15", tga_file(2, tw, th, 15, px16, 0x00))
tga_case("t2_24_id", tga_file(2, tw, th, 24, bgr, 0x20, idlen=7))
tga_case("t3_8", tga_file(3, tw, th, 8, grey, 0x20))
tga_case("t3_16", tga_file(3, tw, th, 16, ga, 0x00))
tga_case("t10_24", tga_file(10, tw, th, 24, tga_rle(bgr, 3, trng), 0x20))
tga_case("t10_32", tga_file(10, tw, th, 32, tga_rle(bgra, 4, trng), 0x08))
tga_case("t10_16", tga_file(10, tw, th, 16, tga_rle(px16, 2, trng), 0x20))
tga_case("t11_8", tga_file(11, tw, th, 8, tga_rle(grey, 1, trng), 0x00))
tga_case("t11_16", tga_file(11, tw, th, 16, tga_rle(ga, 2, trng), 0x20))
pal_n = 40
idx8 = (trng.integers(0, pal_n + 8, tw * th)).astype(np.uint8)  # some indices past the palette -> entry 0
idx8.reshape(th, tw)[:, :5] = idx8.reshape(th, tw)[:, :1]
idx16 = idx8.astype("<u2").tobytes()
pal24 = trng.integers(0, 256, pal_n * 3, dtype=np.uint8).tobytes()
pal32 = trng.integers(0, 256, pal_n * 4, dtype=np.uint8).tobytes()
pal16 = trng.integers(0, 65536, pal_n, dtype=np.uint16).astype("<u2").tobytes()
pal8 = trng.integers(0, 256, pal_n, dtype=np.uint8).tobytes()
tga_case("t1_pal24", tga_file(1, tw, th, 8, idx8.tobytes(), 0x20, cmap=pal24, cmap_bits=24))
tga_case("t1_pal32_bl", tga_file(1, tw, th, 8, idx8.tobytes(), 0x00, cmap=pal32, cmap_bits=32))
tga_case("t1_pal16", tga_file(1, tw, th, 8, idx8.tobytes(), 0x20, cmap=pal16, cmap_bits=16))
tga_case("t1_pal15", tga_file(1, tw, th, 8, idx8.tobytes(), 0x20, cmap=pal16, cmap_bits=15))
tga_case("t1_pal8", tga_file(1, tw, th, 8, idx8.tobytes(), 0x20, cmap=pal8, cmap_bits=8))
tga_case("t1_pal24_idx16", tga_file(1, tw, th, 16, idx16, 0x20, cmap=pal24, cmap_bits=24))
tga_case("t1_pal24_first3", tga_file(1, tw, th, 8, idx8.tobytes(), 0x20, cmap=pal24, cmap_bits=24, cmap_first=3))
tga_case("t9_pal24", tga_file(9, tw, th, 8, tga_rle(idx8.tobytes(), 1, trng), 0x20, cmap=pal24, cmap_bits=24))
tga_case("t9_pal32_idx16", tga_file(9, tw, th, 16, tga_rle(idx16, 2, trng), 0x00, cmap=pal32, cmap_bits=32))
np.savez_compressed(os.path.join(HERE, "stb_image_tga.npz"), **tout)
print("stb_image_tga:", len(tout) // 2, "files")
