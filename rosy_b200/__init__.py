"""rosy_b200 - B200-native drop-in for the texture stage of btipling/Rosy's asset Packager.

The product is the C-ABI shared library `rosy_b200/csrc/librosy_tex_cuda.so` (hand-written sm_100a
CUDA kernels behind include/rosy_tex_cuda.h).  This package is the thin host-side mirror used by
tests/ and bench.py: same function names and error behaviour as the reference's
src/Packager/Packager.h:8-9.  It never falls back to a CPU path: if the extension is missing the
import fails loudly, and without an sm_100 device every compute call raises RosyTexError.
"""
from .texture import (  # noqa: F401
    FORMAT_BC4,
    FORMAT_BC5,
    FORMAT_BC7,
    KIND_COLOR,
    KIND_METALLIC_ROUGHNESS,
    KIND_MIXMAP,
    KIND_NORMAL_MAP,
    NORMALIZE_OFF,
    NORMALIZE_UNPACK,
    QUALITY_FAST,
    QUALITY_HIGH,
    QUALITY_NORMAL,
    LIB_PATH,
    RosyTexError,
    build,
    lib,
)
from . import texture  # noqa: F401
