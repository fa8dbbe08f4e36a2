"""Deterministic synthetic textures for the Packager texture stage.

There is no network, so no real Sponza: every input used by tests/ and bench.py is generated here
from an integer hash (seed 0x524F5359 + texture index).  The three kinds mirror the three glTF
slots the reference routes through the stage (src/Packager/Gltf.cpp:225, :252, :280):

  albedo           5-octave value noise per channel + 64-px checker tint + hard edges; sRGB;
                   optional cut-out alpha mask (about 30 % transparent texels)
  normal           6-octave fBm height -> central-difference gradient -> unit normal -> RGB8
  metal_rough      R = 255, G = 4-octave noise (roughness), B = thresholded low-frequency mask
  random_rgba      uniform noise: the worst case for BC7
"""
import numpy as np

SEED_BASE = 0x524F5359


def _hash2(ix, iy, seed):
    """32-bit integer hash of lattice coordinates -> float32 in [0,1)."""
    x = ix.astype(np.uint64) & 0xFFFFFFFF
    y = iy.astype(np.uint64) & 0xFFFFFFFF
    h = (x * np.uint64(0x9E3779B1) ^ y * np.uint64(0x85EBCA77) ^ np.uint64(seed & 0xFFFFFFFF)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    return (h >> np.uint64(8)).astype(np.float32) * np.float32(1.0 / (1 << 24))


def value_noise(w, h, cell, seed):
    """Smooth value noise with lattice spacing `cell` pixels, float32 (h, w) in [0,1)."""
    cell = max(1, int(cell))
    gx = np.arange(w, dtype=np.float32) / cell
    gy = np.arange(h, dtype=np.float32) / cell
    x0 = np.floor(gx).astype(np.int64)
    y0 = np.floor(gy).astype(np.int64)
    fx = gx - x0
    fy = gy - y0
    fx = fx * fx * (3 - 2 * fx)
    fy = fy * fy * (3 - 2 * fy)
    lat = _hash2(
        np.arange(x0.max() + 2, dtype=np.int64)[None, :], np.arange(y0.max() + 2, dtype=np.int64)[:, None], seed
    )
    a = lat[y0][:, x0]
    b = lat[y0][:, x0 + 1]
    c = lat[y0 + 1][:, x0]
    d = lat[y0 + 1][:, x0 + 1]
    top = a + (b - a) * fx[None, :]
    bot = c + (d - c) * fx[None, :]
    return (top + (bot - top) * fy[:, None]).astype(np.float32)


def fbm(w, h, octaves, seed, base_cell=None):
    base_cell = base_cell or max(4, min(w, h) // 4)
    out = np.zeros((h, w), dtype=np.float32)
    amp, total, cell = 1.0, 0.0, float(base_cell)
    for o in range(octaves):
        out += amp * value_noise(w, h, max(1, int(round(cell))), seed * 131 + o * 7919 + 17)
        total += amp
        amp *= 0.5
        cell *= 0.5
    return out / total


def _to_u8(x):
    return np.clip(np.floor(x * 255.0 + 0.5), 0, 255).astype(np.uint8)


def albedo(w, h, index=0, cutout=None):
    """sRGB albedo.  cutout=None: every 8th texture index gets an alpha mask."""
    seed = SEED_BASE + index
    img = np.empty((h, w, 4), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    checker = (((xx // 64) + (yy // 64)) & 1).astype(np.float32)
    edges = value_noise(w, h, max(8, min(w, h) // 16), seed * 3 + 5) > 0.78  # hard-edged patches
    for c in range(3):
        v = fbm(w, h, 5, seed * 4 + c, base_cell=64)
        v = v + 0.04 * (_hash2(xx, yy, seed * 13 + c) - 0.5)  # sensor-like grain
        v = 0.15 + 0.7 * v + 0.08 * (checker - 0.5) * (1 if c != 1 else -1)
        v = np.where(edges, 1.0 - v, v)
        img[..., c] = _to_u8(v)
    img[..., 3] = 255
    if cutout is None:
        cutout = index % 8 == 7
    if cutout:
        m = fbm(w, h, 3, seed * 5 + 11, base_cell=max(8, min(w, h) // 8))
        thr = np.quantile(m, 0.30)
        soft = np.clip((m - thr) * 40.0 + 0.5, 0.0, 1.0)  # thin anti-aliased rim, ~30 % at 0
        img[..., 3] = _to_u8(soft)
    return img


def normal(w, h, index=0, strength=2.0):
    seed = SEED_BASE + index
    hgt = fbm(w, h, 6, seed * 6 + 1) * (min(w, h) / 16.0)
    dx = (np.roll(hgt, -1, axis=1) - np.roll(hgt, 1, axis=1)) * 0.5 * strength
    dy = (np.roll(hgt, -1, axis=0) - np.roll(hgt, 1, axis=0)) * 0.5 * strength
    n = np.stack([-dx, -dy, np.ones_like(dx)], axis=-1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    img = np.empty((h, w, 4), dtype=np.uint8)
    img[..., :3] = _to_u8(n * 0.5 + 0.5)
    img[..., 3] = 255
    return img


def metal_rough(w, h, index=0):
    seed = SEED_BASE + index
    img = np.empty((h, w, 4), dtype=np.uint8)
    img[..., 0] = 255
    img[..., 1] = _to_u8(0.1 + 0.8 * fbm(w, h, 4, seed * 7 + 2))
    m = value_noise(w, h, max(16, min(w, h) // 8), seed * 9 + 3)
    img[..., 2] = np.where(m > np.quantile(m, 0.8), 255, 0).astype(np.uint8)
    img[..., 3] = 255
    return img


def random_rgba(w, h, index=0):
    rng = np.random.default_rng(SEED_BASE + index)
    return rng.integers(0, 256, (h, w, 4), dtype=np.uint8)


def constant(w, h, rgba):
    img = np.empty((h, w, 4), dtype=np.uint8)
    img[...] = np.asarray(rgba, dtype=np.uint8)
    return img


def gradient(w, h):
    img = np.empty((h, w, 4), dtype=np.uint8)
    img[..., 0] = (np.arange(w) * 255 // max(1, w - 1))[None, :]
    img[..., 1] = (np.arange(h) * 255 // max(1, h - 1))[:, None]
    img[..., 2] = 128
    img[..., 3] = 255
    return img


KINDS = {"albedo": albedo, "normal": normal, "metal_rough": metal_rough, "random": random_rgba}


def make(kind, w, h, index=0):
    return KINDS[kind](w, h, index)
