"""cs378hgraphics_raytracer_b200 - ctypes binding of librt_b200.so (the C-ABI in include/rt_b200.h).

The product is the CUDA library; this module only loads it and wraps the C entry points for the
tests and bench.py.  There is no Python or CPU implementation of the trace loop here: if the
library is missing or no CUDA device is usable, the calls raise.
"""
from .binding import (  # noqa: F401
    LIB_PATH, RtError, Params, Stats, Ray, Hit, HIT_DTYPE, lib, load_ray_file, Arena, Context, build_library, write_image, read_image, save_arena, load_arena, MultiContext,
)
