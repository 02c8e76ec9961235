"""ctypes view of include/rt_b200.h.  Mirrors the C structs field for field."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB: alternative build of the same library (kernel tuning experiments)
LIB_PATH = os.environ.get("RT_B200_LIB", os.path.join(HERE, "librt_b200.so"))

RT_AA_NONE, RT_AA_SUPERSAMPLE, RT_AA_ADAPTIVE, RT_AA_JITTERED = 0, 1, 2, 3
_STATUS = {1: "RT_ERR_ARG", 2: "RT_ERR_CUDA", 3: "RT_ERR_NO_DEVICE", 4: "RT_ERR_NO_SCENE", 5: "RT_ERR_IO",
           6: "RT_ERR_PARSE", 7: "RT_ERR_OVERFLOW", 8: "RT_ERR_UNSUPPORTED"}


class RtError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("%s: %s" % (_STATUS.get(status, status), message))
        self.status = status


class Params(ctypes.Structure):
    """rt_params: the reference's traceUI settings read by the hot loop (TraceUI.h:35-129)."""
    _fields_ = [(n, ctypes.c_int32) for n in
                "depth aa_mode aa_samples dof_on dof_divs anaglyph overlapping_objects use_cubemap soft_shadow_res pad".split()] + \
               [(n, ctypes.c_double) for n in "aterm_thresh aa_thresh dof_fd dof_apsz".split()]

    @classmethod
    def default(cls, **kw):
        p = cls()
        lib().rt_params_default(ctypes.byref(p))
        for k, v in kw.items():
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
        return p


class Stats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint64) for n in "camera_rays reflection_rays refraction_rays shadow_rays".split()] + \
               [("device_ms", ctypes.c_double), ("launches", ctypes.c_uint64), ("cooperative_rays", ctypes.c_uint64),
                ("rounds", ctypes.c_uint64), ("stage_ms", ctypes.c_double * 6), ("stage_launches", ctypes.c_uint64 * 6),
                ("culled_shadow_rays", ctypes.c_uint64)]

    STAGES = ("gen", "closest", "coop", "shade", "shadow", "other")

    @property
    def rays(self):
        return self.camera_rays + self.reflection_rays + self.refraction_rays + self.shadow_rays


class Ray(ctypes.Structure):
    _fields_ = [("p", ctypes.c_double * 3), ("d", ctypes.c_double * 3)]


class Hit(ctypes.Structure):
    _fields_ = [("obj", ctypes.c_int32), ("face", ctypes.c_int32), ("t", ctypes.c_double), ("N", ctypes.c_double * 3)]


HIT_DTYPE = np.dtype([("obj", "i4"), ("face", "i4"), ("t", "f8"), ("N", "f8", 3)])

_lib = None


def build_library():
    """Compiles librt_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", os.path.join(HERE, "csrc"), "-j8"])


def lib():
    """Loads the CUDA library.  Fails loudly: there is no fallback implementation."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtError(3, "librt_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                         "there is no CPU fallback")
    L = ctypes.CDLL(LIB_PATH)
    vp, sz, i32, cp = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int32, ctypes.c_char_p
    L.rt_params_default.argtypes = [ctypes.POINTER(Params)]
    L.rt_params_default.restype = None
    L.rt_create.argtypes = [ctypes.c_int, ctypes.POINTER(vp)]
    L.rt_destroy.argtypes = [vp]
    L.rt_destroy.restype = None
    L.rt_last_error.argtypes = [vp]
    L.rt_last_error.restype = cp
    L.rt_upload_scene.argtypes = [vp, vp, sz]
    L.rt_render.argtypes = [vp, ctypes.POINTER(Params), ctypes.c_int, ctypes.c_int, vp, ctypes.POINTER(Stats)]
    L.rt_render_tiles.argtypes = [vp, ctypes.POINTER(Params), ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp,
                                  ctypes.POINTER(Stats)]
    L.rt_render_device.argtypes = [vp, ctypes.POINTER(Params), ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                   ctypes.POINTER(vp), ctypes.POINTER(Stats)]
    L.rt_trace_pixels.argtypes = [vp, ctypes.POINTER(Params), ctypes.c_int, ctypes.c_int, vp, ctypes.c_int, vp]
    L.rt_trace_rays.argtypes = [vp, ctypes.POINTER(Params), vp, ctypes.c_int, ctypes.c_int, vp, vp]
    L.rt_multi_create.argtypes = [ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.POINTER(vp)]
    L.rt_multi_create.restype = i32
    L.rt_multi_destroy.argtypes = [vp]
    L.rt_multi_destroy.restype = None
    L.rt_multi_last_error.argtypes = [vp]
    L.rt_multi_last_error.restype = cp
    L.rt_multi_devices.argtypes = [vp]
    L.rt_multi_upload_scene.argtypes = [vp, vp, sz]
    L.rt_multi_upload_scene.restype = i32
    L.rt_multi_render.argtypes = [vp, ctypes.POINTER(Params), ctypes.c_int, ctypes.c_int, vp, ctypes.POINTER(Stats)]
    L.rt_multi_render.restype = i32
    L.rt_scene_buffer.argtypes = [vp, sz, ctypes.POINTER(vp)]
    L.rt_scene_buffer.restype = i32
    L.rt_upload_scene_resident.argtypes = [vp, vp, sz]
    L.rt_upload_scene_resident.restype = i32
    L.rt_pack_tiles.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp]
    L.rt_pack_tiles.restype = i32
    L.rt_unpack_tiles.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, vp]
    L.rt_unpack_tiles.restype = i32
    L.rt_intersect_static.argtypes = [vp, vp, ctypes.c_int, vp]
    L.rt_intersect_static.restype = i32
    L.rt_selftest_division.argtypes = [vp, ctypes.c_longlong, ctypes.c_ulonglong, ctypes.c_int, ctypes.POINTER(ctypes.c_ulonglong)]
    L.rt_selftest_division.restype = i32
    L.rt_selftest_derive.argtypes = [vp, vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_ulonglong)]
    L.rt_selftest_derive.restype = i32
    L.rt_measure_fma_peak.argtypes = [vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
    L.rt_measure_fma_peak.restype = i32
    L.rt_set_profiling.argtypes = [vp, ctypes.c_int]
    L.rt_set_profiling.restype = i32
    L.rt_load_ray_file.argtypes = [cp, vp, ctypes.POINTER(vp), ctypes.POINTER(sz)]
    L.rt_arena_free.argtypes = [vp]
    L.rt_arena_free.restype = None
    L.rt_host_last_error.restype = cp
    L.rt_arena_validate.argtypes = [vp, sz]
    L.rt_arena_save.argtypes = [cp, vp, sz]
    L.rt_arena_save.restype = i32
    L.rt_arena_load.argtypes = [cp, ctypes.POINTER(vp), ctypes.POINTER(sz)]
    L.rt_arena_load.restype = i32
    L.rt_write_image.argtypes = [cp, ctypes.c_int, ctypes.c_int, vp]
    L.rt_write_image.restype = i32
    L.rt_read_image.argtypes = [cp, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.POINTER(vp)]
    L.rt_read_image.restype = i32
    for f in ("rt_create", "rt_upload_scene", "rt_render", "rt_render_tiles", "rt_render_device", "rt_trace_pixels",
              "rt_trace_rays", "rt_load_ray_file", "rt_arena_validate"):
        getattr(L, f).restype = i32
    _lib = L
    return L


def _cubemap_arg(faces):
    if faces is None:
        return None
    if len(faces) != 6:
        raise ValueError("cubemap needs six face files (+x -x +y -y +z -z)")
    arr = (ctypes.c_char_p * 6)(*[os.fsencode(f) for f in faces])
    return arr


class Arena:
    """A flattened scene (include/rt_arena.h) in host memory."""

    def __init__(self, ptr, nbytes, owner="product"):
        self.ptr, self.nbytes, self._owner = ptr, nbytes, owner

    def bytes(self):
        return ctypes.string_at(self.ptr, self.nbytes)

    def array(self):
        return np.ctypeslib.as_array(ctypes.cast(self.ptr, ctypes.POINTER(ctypes.c_uint8)), (self.nbytes,))

    def section(self, sec_id, dtype):
        hdr = np.frombuffer(self.array()[:32 + 17 * 24].tobytes(), dtype=np.uint64)
        off, cnt = int(hdr[4 + 3 * sec_id + 1]), int(hdr[4 + 3 * sec_id + 2])
        esz = int(hdr[4 + 3 * sec_id]) >> 32
        return np.frombuffer(self.array()[off:off + cnt * esz].tobytes(), dtype=dtype)

    def camera(self):
        c = self.section(1, np.float64)
        return dict(eye=c[0:3], look=c[3:6], u=c[6:9], v=c[9:12], aspect=float(c[12]))

    def free(self):
        if self.ptr and self._owner == "product":
            lib().rt_arena_free(self.ptr)
        self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def load_ray_file(path, cubemap_faces=None):
    """rt_load_ray_file: parse a .ray scene on the host, build the reference's trees, flatten."""
    L = lib()
    p, n = ctypes.c_void_p(), ctypes.c_size_t()
    cm = _cubemap_arg(cubemap_faces)
    st = L.rt_load_ray_file(os.fsencode(path), cm, ctypes.byref(p), ctypes.byref(n))
    if st != 0:
        raise RtError(st, L.rt_host_last_error().decode(errors="replace"))
    return Arena(p.value, n.value)


def save_arena(path, arena):
    """Writes the flattened scene to a cache file (rt_arena_save)."""
    st = lib().rt_arena_save(os.fsencode(path), ctypes.c_void_p(arena.ptr), arena.nbytes)
    if st != 0:
        raise RtError(st, "could not save the arena to " + str(path))


def load_arena(path):
    """Reads a cache file written by save_arena (rt_arena_load: validated, refused if stale or damaged)."""
    p, n = ctypes.c_void_p(), ctypes.c_size_t()
    st = lib().rt_arena_load(os.fsencode(path), ctypes.byref(p), ctypes.byref(n))
    if st != 0:
        raise RtError(st, "could not load an arena from " + str(path))
    return Arena(p.value, n.value)


def write_image(path, img):
    """writeImage (images.cc:58-66): img is (h, w, 3) uint8 with row 0 = bottom scan line; .png or .bmp by extension."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w, _ = img.shape
    st = lib().rt_write_image(os.fsencode(path), w, h, img.ctypes.data_as(ctypes.c_void_p))
    if st != 0:
        raise RtError(st, "could not write " + str(path))


def read_image(path):
    """readImage (images.cc:48-55): (h, w, 3) uint8, row 0 = bottom scan line."""
    w, h, p = ctypes.c_int(), ctypes.c_int(), ctypes.c_void_p()
    st = lib().rt_read_image(os.fsencode(path), ctypes.byref(w), ctypes.byref(h), ctypes.byref(p))
    if st != 0:
        raise RtError(st, "could not read " + str(path))
    try:
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_uint8)), (h.value, w.value, 3)).copy()
    finally:
        lib().rt_arena_free(p)


class Context:
    """rt_ctx: one GPU.  Mirrors the part of RayTracer the C-ABI replaces."""

    def __init__(self, device=0):
        L = lib()
        h = ctypes.c_void_p()
        st = L.rt_create(device, ctypes.byref(h))
        if st != 0:
            raise RtError(st, "rt_create(device=%d) failed: no usable CUDA device; there is no CPU path" % device)
        self._h = h

    def _check(self, st):
        if st != 0:
            raise RtError(st, lib().rt_last_error(self._h).decode(errors="replace"))

    def close(self):
        if getattr(self, "_h", None):
            lib().rt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, arena):
        self._check(lib().rt_upload_scene(self._h, arena.ptr, arena.nbytes))

    def render(self, params, w, h, part=0, n_parts=1, out=None):
        """rt_render / rt_render_tiles -> (rgb[h,w,3] uint8 with row 0 = bottom scan line, Stats)."""
        if out is None:
            out = np.zeros((h, w, 3), np.uint8)
        st = Stats()
        self._check(lib().rt_render_tiles(self._h, ctypes.byref(params), w, h, part, n_parts, out.ctypes.data, ctypes.byref(st)))
        return out, st

    def render_device(self, params, w, h, part=0, n_parts=1):
        """rt_render_device -> (device pointer of the RGB8 frame, Stats): nothing crosses PCIe."""
        st = Stats()
        dev = ctypes.c_void_p()
        self._check(lib().rt_render_device(self._h, ctypes.byref(params), w, h, part, n_parts, ctypes.byref(dev), ctypes.byref(st)))
        return dev.value, st

    def selftest_derive(self, arena):
        bad = ctypes.c_ulonglong()
        self._check(lib().rt_selftest_derive(self._h, arena.ptr, arena.nbytes, ctypes.byref(bad)))
        return bad.value

    def measure_fma_peak(self):
        """(FP64, FP32) FMA TFLOP/s of this GPU."""
        a, b = ctypes.c_double(), ctypes.c_double()
        self._check(lib().rt_measure_fma_peak(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def set_profiling(self, on):
        """rt_set_profiling: per-stage CUDA-event times in Stats.stage_ms."""
        self._check(lib().rt_set_profiling(self._h, 1 if on else 0))

    def selftest_division(self, n, seed=1, exp_range=40):
        bad = ctypes.c_ulonglong()
        self._check(lib().rt_selftest_division(self._h, n, seed, exp_range, ctypes.byref(bad)))
        return bad.value

    def trace_pixels(self, params, w, h, ij):
        ij = np.ascontiguousarray(ij, dtype=np.int32).reshape(-1, 2)
        out = np.zeros((len(ij), 3))
        self._check(lib().rt_trace_pixels(self._h, ctypes.byref(params), w, h, ij.ctypes.data, len(ij), out.ctypes.data))
        return out

    def pack_tiles(self, dev_rgb, w, h, part, n_parts, dev_packed):
        """rt_pack_tiles: this partition's tiles of the device frame -> contiguous device block (device pointers as ints)."""
        self._check(lib().rt_pack_tiles(self._h, ctypes.c_void_p(dev_rgb), w, h, part, n_parts, ctypes.c_void_p(dev_packed)))

    def unpack_tiles(self, dev_gathered, w, h, n_parts, tiles_cap, dev_rgb_out):
        self._check(lib().rt_unpack_tiles(self._h, ctypes.c_void_p(dev_gathered), w, h, n_parts, tiles_cap, ctypes.c_void_p(dev_rgb_out)))

    def intersect_static(self, rays):
        """rt_intersect_static: Scene::intersect for at most STATIC_MAX_RAYS rays, no device allocation."""
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
        hits = np.zeros(len(rays), HIT_DTYPE)
        self._check(lib().rt_intersect_static(self._h, rays.ctypes.data, len(rays), hits.ctypes.data))
        return hits

    def trace_rays(self, params, rays, depth, want_hits=True, want_rgb=True):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
        n = len(rays)
        hits = np.zeros(n, HIT_DTYPE) if want_hits else None
        rgb = np.zeros((n, 3)) if want_rgb else None
        self._check(lib().rt_trace_rays(self._h, ctypes.byref(params), rays.ctypes.data, n, depth,
                                        hits.ctypes.data if want_hits else None, rgb.ctypes.data if want_rgb else None))
        return hits, rgb


class MultiContext:
    """rt_multi_*: N GPUs of one box in ONE process (scene replicated by NCCL broadcast, interleaved tiles, gather to
    the first device).  No fallback: fails loudly without the devices, or - for N > 1 - without NCCL."""

    def __init__(self, devices):
        arr = (ctypes.c_int * len(devices))(*devices)
        h = ctypes.c_void_p()
        st = lib().rt_multi_create(arr, len(devices), ctypes.byref(h))
        if st != 0:
            raise RtError(st, "rt_multi_create(%s) failed" % list(devices))
        self._h = h

    def _check(self, st):
        if st != 0:
            raise RtError(st, lib().rt_multi_last_error(self._h).decode(errors="replace"))

    def upload(self, arena):
        self._check(lib().rt_multi_upload_scene(self._h, arena.ptr, arena.nbytes))

    def render(self, params, w, h):
        out = np.zeros((h, w, 3), np.uint8)
        st = Stats()
        self._check(lib().rt_multi_render(self._h, ctypes.byref(params), w, h, out.ctypes.data, ctypes.byref(st)))
        return out, st

    def close(self):
        if self._h:
            lib().rt_multi_destroy(self._h)
            self._h = None
