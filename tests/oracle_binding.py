"""ctypes view of the oracle: oracle/_ref/libref_oracle.so = the UNMODIFIED reference ray tracer
(compiled from /root/reference where it exists, shipped prebuilt to the GPU box) behind the
extern "C" harness oracle/ref_harness.cpp.  Test infrastructure only."""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_oracle.so")

_ref = None


def available():
    return os.path.exists(REF_LIB)


def ref():
    global _ref
    if _ref is None:
        L = ctypes.CDLL(REF_LIB)
        L.ref_load.argtypes = [ctypes.c_char_p, ctypes.c_void_p]
        L.ref_load_arena.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
        L.ref_flatten.restype = ctypes.c_void_p
        L.ref_flatten.argtypes = [ctypes.POINTER(ctypes.c_size_t)]
        L.ref_free.argtypes = [ctypes.c_void_p]
        L.ref_aspect.restype = ctypes.c_double
        L.ref_last_alert.restype = ctypes.c_char_p
        L.ref_set_params.argtypes = [ctypes.c_void_p]
        L.ref_set_threads.argtypes = [ctypes.c_int]
        L.ref_trace_rays.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        L.ref_trace_pixels.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.ref_render.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.ref_render_window.argtypes = [ctypes.c_int] * 6 + [ctypes.c_void_p] * 3
        _ref = L
    return _ref


class Oracle:
    """The reference RayTracer with one scene loaded (the harness keeps a single global scene,
    like the reference's own globals)."""

    def __init__(self, path, cubemap_faces=None, arena=None):
        from cs378hgraphics_raytracer_b200.binding import _cubemap_arg
        L = ref()
        if arena is not None:
            # populate the reference's own classes from an arena (ref_harness.cpp: ref_load_arena);
            # used for scenes whose reference-side load would take tens of minutes
            rc = L.ref_load_arena(arena.ptr, arena.nbytes)
            if rc != 0:
                raise RuntimeError("ref_load_arena failed: %d" % rc)
            self.aspect = L.ref_aspect()
            return
        self._cm = _cubemap_arg(cubemap_faces)
        if L.ref_load(os.fsencode(path), self._cm) != 0:
            raise RuntimeError("reference failed to load %s: %s" % (path, L.ref_last_alert().decode(errors="replace")))
        self.aspect = L.ref_aspect()

    def set_params(self, params):
        ref().ref_set_params(ctypes.byref(params))

    def flatten(self):
        n = ctypes.c_size_t()
        p = ref().ref_flatten(ctypes.byref(n))
        data = ctypes.string_at(p, n.value)
        ref().ref_free(p)
        return data

    def trace_rays(self, params, rays, depth, want_hits=True, want_rgb=True):
        from cs378hgraphics_raytracer_b200.binding import HIT_DTYPE
        self.set_params(params)
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
        n = len(rays)
        hits = np.zeros(n, HIT_DTYPE) if want_hits else None
        rgb = np.zeros((n, 3)) if want_rgb else None
        rc = ref().ref_trace_rays(rays.ctypes.data, n, depth, hits.ctypes.data if want_hits else None,
                                  rgb.ctypes.data if want_rgb else None)
        assert rc == 0
        return hits, rgb

    def trace_pixels(self, params, w, h, ij):
        self.set_params(params)
        ij = np.ascontiguousarray(ij, dtype=np.int32).reshape(-1, 2)
        out = np.zeros((len(ij), 3))
        assert ref().ref_trace_pixels(w, h, ij.ctypes.data, len(ij), out.ctypes.data) == 0
        return out

    def render(self, params, w, h):
        from cs378hgraphics_raytracer_b200.binding import Stats
        assert (w, h) != (256, 256), "the reference writes through a null buffer at exactly 256x256 (SURVEY A.9)"
        self.set_params(params)
        out = np.zeros((h, w, 3), np.uint8)
        sec = ctypes.c_double()
        st = Stats()
        assert ref().ref_render(w, h, out.ctypes.data, ctypes.byref(sec), ctypes.byref(st)) == 0
        return out, st, sec.value

    def render_window(self, params, w, h, i0, i1, j0, j1):
        from cs378hgraphics_raytracer_b200.binding import Stats
        self.set_params(params)
        out = np.zeros((j1 - j0, i1 - i0, 3), np.uint8)
        sec = ctypes.c_double()
        st = Stats()
        assert ref().ref_render_window(w, h, i0, i1, j0, j1, out.ctypes.data, ctypes.byref(sec), ctypes.byref(st)) == 0
        return out, st, sec.value


def frame_height(w, aspect):
    """CommandLineUI::run (reference ui/CommandLineUI.cpp:155-156)."""
    return int(w / aspect + 0.5)


def camera_rays(cam, w, h):
    """Camera::rayThrough for every pixel (i,j) of a w x h frame at x=i/w, y=j/h
    (reference scene/camera.cpp:20-30, RayTracer.cpp:87-88), in glm's evaluation order."""
    i, j = np.meshgrid(np.arange(w), np.arange(h), indexing="xy")
    x = (i.reshape(-1) / float(w)) - 0.5
    y = (j.reshape(-1) / float(h)) - 0.5
    d = (cam["look"][None, :] + x[:, None] * cam["u"][None, :]) + y[:, None] * cam["v"][None, :]
    inv = 1.0 / np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2])
    d = d * inv[:, None]
    rays = np.zeros((w * h, 6))
    rays[:, 0:3] = cam["eye"]
    rays[:, 3:6] = d
    return rays


def image_diff(a, b):
    """raycheck.py-style comparison (reference ray/raycheck.py:21-34) done in integers (the script's
    uint8 subtraction wraps), plus the north-star metric: share of pixels whose worst channel
    differs by at most 1."""
    d = np.abs(a.astype(np.int32) - b.astype(np.int32))
    worst = d.max(axis=2)
    return dict(rms=float(np.sqrt((d.astype(np.float64) ** 2).mean())), within1=float((worst <= 1).mean()),
                exact=float((worst == 0).mean()), max=int(worst.max()))
