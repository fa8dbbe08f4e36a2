"""ctypes mirror of include/rosy_tex_cuda.h (the reference-facing boundary).

`generate_srgb_texture(path)` / `generate_normal_map_texture(path)` mirror
/root/reference/src/Packager/Packager.h:8-9: they read the image, write `<path>.dds` next to it and
raise RosyTexError where the reference returns rosy::result::error.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("ROSY_TEX_LIB") or os.path.join(CSRC, "librosy_tex_cuda.so")  # override: A/B builds only

KIND_COLOR, KIND_NORMAL_MAP, KIND_METALLIC_ROUGHNESS, KIND_MIXMAP = 0, 1, 2, 3
FORMAT_BC7, FORMAT_BC5, FORMAT_BC4 = 0, 1, 2
QUALITY_FAST, QUALITY_NORMAL, QUALITY_HIGH = 0, 1, 2
NORMALIZE_OFF, NORMALIZE_UNPACK = 0, 1
DDS_HEADER_BYTES = 148
STATUS = {0: "OK", 1: "IO", 2: "DECODE", 3: "ARG", 4: "CUDA", 5: "NO_DEVICE", 6: "NOMEM"}


class RosyTexError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"rosy_tex status {status} ({STATUS.get(status, '?')}): {message}")
        self.status = status


class Desc(C.Structure):
    _fields_ = [
        ("width", C.c_uint32),
        ("height", C.c_uint32),
        ("kind", C.c_int32),
        ("format", C.c_int32),
        ("quality", C.c_int32),
        ("normalize", C.c_int32),
    ]


class BatchItem(C.Structure):
    _fields_ = [
        ("desc", Desc),
        ("d_rgba", C.c_void_p),
        ("d_chain", C.c_void_p),
        ("d_scratch", C.c_void_p),
        ("d_blocks", C.c_void_p),
    ]


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("num_devices", C.c_int32),
        ("devices", C.c_int32 * 16),
        ("quality", C.c_int32),
        ("format_for_kind", C.c_int32 * 4),
        ("normalize", C.c_int32),
        ("streams_per_device", C.c_int32),
        ("linear_data_maps", C.c_int32),
        ("srgb_flag", C.c_int32),
    ]


def build(verbose=False):
    """Compile the CUDA extension in tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-s", "-C", CSRC], stdout=out)
    return LIB_PATH


_lib = None


def lib():
    """The loaded C-ABI library.  Fails loudly when it has not been built: no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C {CSRC}` "
                "(or __graft_entry__.build()); there is no CPU fallback"
            )
        L = C.CDLL(LIB_PATH)
        u8p = C.POINTER(C.c_uint8)
        L.rosy_tex_last_error.restype = C.c_char_p
        L.rosy_tex_status_string.restype = C.c_char_p
        L.rosy_tex_dds_size.restype = C.c_size_t
        L.rosy_tex_chain_size.restype = C.c_size_t
        L.rosy_tex_device_scratch_size.restype = C.c_size_t
        L.rosy_tex_kernel_launches.restype = C.c_uint64
        L.rosy_tex_debug_faults.restype = C.c_longlong
        L.rosy_tex_alloc_host.restype = C.c_void_p
        L.rosy_tex_alloc_host.argtypes = [C.c_size_t]
        L.rosy_tex_free_host.argtypes = [C.c_void_p]
        L.rosy_tex_chain_size.argtypes = [C.c_uint32, C.c_uint32]
        L.rosy_tex_count_mipmaps.argtypes = [C.c_uint32, C.c_uint32]
        for name in ("rosy_tex_encode_rgba8", "rosy_tex_mips_rgba8", "rosy_tex_compress_level_rgba8",
                     "rosy_tex_submit_rgba8"):
            getattr(L, name).argtypes = [C.POINTER(Desc), C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.rosy_tex_mips_device.argtypes = [C.POINTER(Desc)] + [C.c_void_p] * 4
        L.rosy_tex_mips_device_static.argtypes = [C.POINTER(Desc)] + [C.c_void_p] * 3
        L.rosy_tex_blocks_device.argtypes = [C.POINTER(Desc)] + [C.c_void_p] * 4
        L.rosy_tex_encode_device.argtypes = [C.POINTER(Desc)] + [C.c_void_p] * 5
        L.rosy_tex_batch_scratch_size.restype = C.c_size_t
        L.rosy_tex_batch_scratch_size.argtypes = [C.c_int]
        L.rosy_tex_encode_batch_device.argtypes = [C.POINTER(BatchItem), C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.rosy_tex_load_image.argtypes = [C.c_char_p, C.POINTER(u8p), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.rosy_tex_free_image.argtypes = [u8p]
        L.rosy_tex_generate_texture_file.argtypes = [C.c_char_p, C.c_int, C.c_char_p]
        L.rosy_tex_submit.argtypes = [C.c_char_p, C.c_int]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RosyTexError(rc, lib().rosy_tex_last_error().decode("utf-8", "replace"))


def desc(width, height, kind=KIND_COLOR, format=-1, quality=-1, normalize=-1):
    return Desc(int(width), int(height), int(kind), int(format), int(quality), int(normalize))


def init(devices=None, quality=QUALITY_NORMAL, normalize=NORMALIZE_UNPACK, formats=None, streams_per_device=2,
         linear_data_maps=False, srgb_flag=False):
    cfg = Config()
    cfg.struct_size = C.sizeof(Config)
    if devices:
        cfg.num_devices = len(devices)
        for i, d in enumerate(devices):
            cfg.devices[i] = d
    cfg.quality = quality
    cfg.normalize = normalize
    cfg.streams_per_device = streams_per_device
    cfg.linear_data_maps = int(linear_data_maps)
    cfg.srgb_flag = int(srgb_flag)
    for k, f in (formats or {}).items():
        cfg.format_for_kind[k] = f
    _check(lib().rosy_tex_init(C.byref(cfg)))


def shutdown():
    lib().rosy_tex_shutdown()


def device_count():
    return lib().rosy_tex_device_count()


# ---- 1:1 with the reference (Packager.h:8-9) ----
def generate_srgb_texture(image_path):
    _check(lib().rosy_tex_generate_srgb_texture(os.fsencode(image_path)))


def generate_normal_map_texture(image_path):
    _check(lib().rosy_tex_generate_normal_map_texture(os.fsencode(image_path)))


def generate_texture_file(image_path, kind, dds_path=None):
    _check(lib().rosy_tex_generate_texture_file(os.fsencode(image_path), int(kind),
                                                os.fsencode(dds_path) if dds_path else None))


def load_image(image_path):
    p = C.POINTER(C.c_uint8)()
    w, h = C.c_uint32(), C.c_uint32()
    _check(lib().rosy_tex_load_image(os.fsencode(image_path), C.byref(p), C.byref(w), C.byref(h)))
    arr = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
    lib().rosy_tex_free_image(p)
    return arr


# ---- geometry ----
def count_mipmaps(w, h):
    return lib().rosy_tex_count_mipmaps(w, h)


def level_extent(w, h, level):
    lw, lh = C.c_uint32(), C.c_uint32()
    lib().rosy_tex_level_extent(C.c_uint32(w), C.c_uint32(h), int(level), C.byref(lw), C.byref(lh))
    return lw.value, lh.value


def dds_size(d):
    return lib().rosy_tex_dds_size(C.byref(d))


def chain_size(w, h):
    return lib().rosy_tex_chain_size(w, h)


def dds_header(d):
    out = np.zeros(DDS_HEADER_BYTES, dtype=np.uint8)
    _check(lib().rosy_tex_dds_header(C.byref(d), out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    assert a.ndim == 3 and a.shape[2] == 4, a.shape
    return a


# ---- memory to memory ----
def encode_rgba8(img, kind=KIND_COLOR, format=-1, quality=-1, normalize=-1):
    """Whole stage for one decoded image -> the bytes of its .dds (numpy uint8)."""
    img = _img(img)
    d = desc(img.shape[1], img.shape[0], kind, format, quality, normalize)
    out = np.empty(dds_size(d), dtype=np.uint8)
    n = C.c_size_t()
    _check(lib().rosy_tex_encode_rgba8(C.byref(d), img.ctypes.data, out.ctypes.data, out.size, C.byref(n)))
    assert n.value == out.size
    return out


def mips_rgba8(img, kind=KIND_COLOR, normalize=-1):
    """RGBA8 mip chain as the encoder sees it -> list of (h, w, 4) arrays, level 0 first."""
    img = _img(img)
    h, w = img.shape[:2]
    d = desc(w, h, kind, -1, -1, normalize)
    out = np.empty(chain_size(w, h), dtype=np.uint8)
    n = C.c_size_t()
    _check(lib().rosy_tex_mips_rgba8(C.byref(d), img.ctypes.data, out.ctypes.data, out.size, C.byref(n)))
    levels, off = [], 0
    for l in range(count_mipmaps(w, h)):
        lw, lh = level_extent(w, h, l)
        levels.append(out[off : off + 4 * lw * lh].reshape(lh, lw, 4))
        off += 4 * lw * lh
    return levels


def compress_level(img, format=FORMAT_BC7, quality=-1):
    """Block-compress one level as is (Context::compress of a single surface) -> raw blocks."""
    img = _img(img)
    h, w = img.shape[:2]
    d = desc(w, h, KIND_COLOR, format, quality, -1)
    nb = ((w + 3) // 4) * ((h + 3) // 4)
    out = np.empty(nb * (8 if format == FORMAT_BC4 else 16), dtype=np.uint8)
    n = C.c_size_t()
    _check(lib().rosy_tex_compress_level_rgba8(C.byref(d), img.ctypes.data, out.ctypes.data, out.size, C.byref(n)))
    assert n.value == out.size
    return out


# ---- asynchronous multi-device work queue ----
class Pending:
    """Keeps the buffers of submitted textures alive until wait_all()."""

    def __init__(self):
        self.items = []

    def submit(self, img, kind=KIND_COLOR, format=-1, quality=-1, normalize=-1):
        img = _img(img)
        d = desc(img.shape[1], img.shape[0], kind, format, quality, normalize)
        out = np.empty(dds_size(d), dtype=np.uint8)
        n = C.c_size_t()
        self.items.append((d, img, out, n))
        _check(lib().rosy_tex_submit_rgba8(C.byref(d), img.ctypes.data, out.ctypes.data, out.size, C.byref(n)))
        return out

    def wait_all(self):
        _check(lib().rosy_tex_wait_all())
        outs = [it[2] for it in self.items]
        self.items = []
        return outs


def submit_file(image_path, kind):
    _check(lib().rosy_tex_submit(os.fsencode(image_path), int(kind)))


def wait_all():
    _check(lib().rosy_tex_wait_all())


# ---- device-resident (torch tensors are only the allocator / stream provider) ----
def device_scratch_size(d):
    return lib().rosy_tex_device_scratch_size(C.byref(d))


def mips_device(d, rgba_ptr, chain_ptr, scratch_ptr, stream=0):
    _check(lib().rosy_tex_mips_device(C.byref(d), rgba_ptr, chain_ptr, scratch_ptr, stream))


STATIC_SCRATCH_BYTES = 5 * 1024 * 1024 + 256  # ROSY_TEX_STATIC_SCRATCH_BYTES


def mips_device_static(d, rgba_ptr, chain_ptr, stream=0):
    """rosy_tex_mips_device_static: the mip chain for callers that must not allocate device memory.  The
    scratch is a fixed array inside the library (textures of at most 1024 x 1024 pixels); calls on one device
    share it, so keep them on one stream."""
    _check(lib().rosy_tex_mips_device_static(C.byref(d), rgba_ptr, chain_ptr, stream))


def blocks_device(d, rgba_ptr, chain_ptr, blocks_ptr, stream=0):
    _check(lib().rosy_tex_blocks_device(C.byref(d), rgba_ptr, chain_ptr, blocks_ptr, stream))


def encode_device(d, rgba_ptr, chain_ptr, scratch_ptr, blocks_ptr, stream=0):
    _check(lib().rosy_tex_encode_device(C.byref(d), rgba_ptr, chain_ptr, scratch_ptr, blocks_ptr, stream))


def batch_scratch_size(count):
    return lib().rosy_tex_batch_scratch_size(int(count))


def make_batch(entries):
    """entries: (Desc, rgba_ptr, chain_ptr, scratch_ptr, blocks_ptr) per texture -> ctypes array"""
    arr = (BatchItem * len(entries))()
    for it, (d, a, b, c, e) in zip(arr, entries):
        it.desc, it.d_rgba, it.d_chain, it.d_scratch, it.d_blocks = d, a, b, c, e
    return arr


def encode_batch_device(batch, batch_scratch_ptr, stream=0, do_mips=True, do_blocks=True):
    _check(lib().rosy_tex_encode_batch_device(batch, len(batch), batch_scratch_ptr, stream, int(do_mips), int(do_blocks)))


def debug_faults():
    """out-of-bounds stores counted by a `make checked` build of the library; -1 for the product build"""
    return lib().rosy_tex_debug_faults()


def kernel_launches():
    return lib().rosy_tex_kernel_launches()


def queue_status():
    """(live_workers, requeued) of the work queue; live_workers is -1 before the first submit"""
    a, b = C.c_int(0), C.c_int(0)
    lib().rosy_tex_queue_status(C.byref(a), C.byref(b))
    return a.value, b.value
