"""ctypes/numpy bindings of oracle/liboracle.so and oracle/_ref/*.so (test infrastructure)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

KIND_SRGB, KIND_NORMAL = 0, 1
FMT_BC7, FMT_BC5, FMT_BC4 = 0, 1, 2
Q_FAST, Q_NORMAL, Q_HIGH = 0, 1, 2
Q_EXHAUSTIVE = 3  # oracle-only quality ceiling (every partition / rotation / p-bit, LS to convergence)
NORMALIZE_OFF, NORMALIZE_UNPACK = 0, 1
DDS_HEADER_BYTES = 148

_u8p = C.POINTER(C.c_uint8)


def build(ref=True):
    """Compile the oracle (and, when /root/reference is present, oracle/_ref)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "CC=gcc", "CXX=g++"])
    if ref and os.path.isdir("/root/reference/src/libs"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref", "CC=gcc", "CXX=g++"])


def _load():
    path = os.path.join(_HERE, "liboracle.so")
    if not os.path.exists(path):
        build(ref=False)
    lib = C.CDLL(path)
    lib.oracle_chain_pixels.restype = C.c_size_t
    lib.oracle_chain_blocks.restype = C.c_size_t
    lib.oracle_dds_size.restype = C.c_size_t
    lib.oracle_generate_texture.restype = C.c_size_t
    return lib


_lib = _load()


def _ptr(a):
    return a.ctypes.data_as(_u8p)


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    assert a.ndim == 3 and a.shape[2] == 4, a.shape
    return a


def count_mipmaps(w, h):
    return _lib.oracle_count_mipmaps(int(w), int(h))


def level_extent(w, h, level):
    lw, lh = C.c_int(), C.c_int()
    _lib.oracle_level_extent(int(w), int(h), int(level), C.byref(lw), C.byref(lh))
    return lw.value, lh.value


def chain_pixels(w, h):
    return _lib.oracle_chain_pixels(int(w), int(h))


def chain_blocks(w, h):
    return _lib.oracle_chain_blocks(int(w), int(h))


def split_levels(flat, w, h):
    """flat RGBA8 chain -> list of (h_l, w_l, 4) arrays"""
    out, off = [], 0
    for l in range(count_mipmaps(w, h)):
        lw, lh = level_extent(w, h, l)
        out.append(flat[off : off + 4 * lw * lh].reshape(lh, lw, 4))
        off += 4 * lw * lh
    return out


def mip_chain(img, kind, normalize_mode=NORMALIZE_UNPACK):
    img = _img(img)
    h, w = img.shape[:2]
    out = np.empty(4 * chain_pixels(w, h), dtype=np.uint8)
    if kind == KIND_NORMAL:
        _lib.oracle_mip_chain_normal(_ptr(img), w, h, int(normalize_mode), _ptr(out))
    else:
        _lib.oracle_mip_chain_srgb(_ptr(img), w, h, _ptr(out))
    return split_levels(out, w, h)


def block_bytes(fmt):
    return 8 if fmt == FMT_BC4 else 16


def encode_image(img, fmt=FMT_BC7, quality=Q_NORMAL, threads=0):
    img = _img(img)
    h, w = img.shape[:2]
    nb = ((w + 3) // 4) * ((h + 3) // 4)
    out = np.empty(nb * block_bytes(fmt), dtype=np.uint8)
    _lib.oracle_encode_image(_ptr(img), w, h, int(fmt), int(quality), _ptr(out), int(threads))
    return out


def decode_image(blocks, w, h, fmt=FMT_BC7):
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty((h, w, 4), dtype=np.uint8)
    _lib.oracle_decode_image(_ptr(blocks), int(w), int(h), int(fmt), _ptr(out))
    return out


def bc7_encode_block(px64, quality=Q_NORMAL):
    px = np.ascontiguousarray(px64, dtype=np.uint8).reshape(64)
    out = np.zeros(16, dtype=np.uint8)
    _lib.oracle_bc7_encode_block(_ptr(px), int(quality), _ptr(out))
    return out


def bc7_decode_block(blk):
    blk = np.ascontiguousarray(blk, dtype=np.uint8).reshape(16)
    out = np.zeros(64, dtype=np.uint8)
    _lib.oracle_bc7_decode_block(_ptr(blk), _ptr(out))
    return out.reshape(16, 4)


def bc7_block_modes(blocks):
    """mode per block (vectorised): index of the lowest set bit of byte 0, -1 if reserved"""
    b0 = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)[:, 0].astype(np.int32)
    mode = np.full(b0.shape, -1, dtype=np.int32)
    for m in range(7, -1, -1):
        mode[(b0 & ((1 << (m + 1)) - 1)) == (1 << m)] = m
    return mode


def bc4_encode_block(src16):
    s = np.ascontiguousarray(src16, dtype=np.uint8).reshape(16)
    out = np.zeros(8, dtype=np.uint8)
    _lib.oracle_bc4_encode_block(_ptr(s), _ptr(out))
    return out


def bc5_encode_block(rg32):
    s = np.ascontiguousarray(rg32, dtype=np.uint8).reshape(32)
    out = np.zeros(16, dtype=np.uint8)
    _lib.oracle_bc5_encode_block(_ptr(s), _ptr(out))
    return out


def dds_header(w, h, levels, fmt):
    out = np.zeros(DDS_HEADER_BYTES, dtype=np.uint8)
    _lib.oracle_dds_header(int(w), int(h), int(levels), int(fmt), _ptr(out))
    return out


def dds_size(w, h, fmt):
    return _lib.oracle_dds_size(int(w), int(h), int(fmt))


def generate_texture(img, kind, fmt=FMT_BC7, quality=Q_NORMAL, normalize_mode=NORMALIZE_UNPACK, threads=0):
    """generate_srgb_texture / generate_normal_map_texture, memory to memory -> DDS bytes"""
    img = _img(img)
    h, w = img.shape[:2]
    out = np.empty(dds_size(w, h, fmt), dtype=np.uint8)
    n = _lib.oracle_generate_texture(
        _ptr(img), w, h, int(kind), int(fmt), int(quality), int(normalize_mode), _ptr(out), int(threads)
    )
    assert n == out.size
    return out


def max_threads():
    return _lib.oracle_max_threads()


def psnr(a, b, channels=4):
    a = np.asarray(a, dtype=np.float64)[..., :channels]
    b = np.asarray(b, dtype=np.float64)[..., :channels]
    mse = np.mean((a - b) ** 2)
    return 99.0 if mse == 0 else 10.0 * np.log10(255.0 * 255.0 / mse)


# ---------------------------------------------------------------- oracle/_ref (the real reference)
def _load_ref(name):
    path = os.path.join(_HERE, "_ref", name)
    return C.CDLL(path) if os.path.exists(path) else None


_stb = _load_ref("libstb_ref.so")
_dds = _load_ref("libdds_ref.so")


def have_ref_stb():
    return _stb is not None


def have_ref_dds():
    return _dds is not None


_asset = _load_ref("libasset_ref.so")


def have_ref_asset():
    return _asset is not None


def ref_asset_write_images(path, images):
    """the reference's own asset::write (src/Asset/Asset.cpp:39-340) for an asset that holds only `images`
    [(image_type, name), ...]; writes the .rsy at `path`"""
    n = len(images)
    types = (C.c_uint32 * max(n, 1))(*[int(t) for t, _ in images])
    names = (C.c_char_p * max(n, 1))(*[nm.encode() for _, nm in images])
    rc = _asset.ref_asset_write_images(str(path).encode(), n, types, names)
    if rc != 0:
        raise RuntimeError("asset::write failed")


def ref_asset_read_images(path, max_images=256, name_stride=1024):
    """the reference's own asset::read (Asset.cpp:409-848) on the .rsy at `path` ->
    ([(image_type, name), ...], (materials, samplers, scenes, nodes, meshes)); None if it rejects the file"""
    types = (C.c_uint32 * max_images)()
    names = C.create_string_buffer(max_images * name_stride)
    other = (C.c_uint64 * 5)()
    n = _asset.ref_asset_read_images(str(path).encode(), max_images, types, names, name_stride, other)
    if n < 0:
        return None
    raw = names.raw
    out = [(int(types[i]), raw[i * name_stride:(i + 1) * name_stride].split(b"\0", 1)[0].decode()) for i in range(min(n, max_images))]
    return out, tuple(int(x) for x in other)


def ref_bc4_block(src16):
    s = np.ascontiguousarray(src16, dtype=np.uint8).reshape(16)
    out = np.zeros(8, dtype=np.uint8)
    _stb.ref_bc4_block(_ptr(out), _ptr(s))
    return out


def ref_bc5_block(rg32):
    s = np.ascontiguousarray(rg32, dtype=np.uint8).reshape(32)
    out = np.zeros(16, dtype=np.uint8)
    _stb.ref_bc5_block(_ptr(out), _ptr(s))
    return out


def ref_image_load(data: bytes):
    w, h = C.c_int(), C.c_int()
    _stb.ref_image_load.restype = C.c_void_p
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    p = _stb.ref_image_load(buf, len(data), C.byref(w), C.byref(h))
    if not p:
        return None
    arr = np.ctypeslib.as_array(C.cast(p, _u8p), shape=(h.value, w.value, 4)).copy()
    _stb.ref_free(C.c_void_p(p))
    return arr


def ref_resize_srgb_box(img):
    img = _img(img)
    h, w = img.shape[:2]
    ow, oh = max(1, w // 2), max(1, h // 2)
    out = np.empty((oh, ow, 4), dtype=np.uint8)
    ok = _stb.ref_resize_srgb_box(_ptr(img), w, h, _ptr(out), ow, oh)
    assert ok
    return out


def ref_dds_read(data):
    """Parse DDS bytes with the reference's dds::readImage.  Returns dict or raises."""
    data = np.ascontiguousarray(np.frombuffer(bytes(data), dtype=np.uint8))
    out = (C.c_uint32 * 40)()
    rc = _dds.ref_dds_read(_ptr(data), C.c_size_t(data.size), out)
    if rc != 0:
        raise ValueError(f"dds::readImage failed: ReadResult {rc - 1}")
    n = out[7]
    return {
        "numMips": out[0],
        "arraySize": out[1],
        "width": out[2],
        "height": out[3],
        "dimension": out[4],
        "format": out[5],
        "supportsAlpha": out[6],
        "mips": [(out[8 + 2 * i], out[9 + 2 * i]) for i in range(min(n, 16))],
    }
