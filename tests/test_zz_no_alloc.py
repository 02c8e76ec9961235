"""rt_intersect_static - the allocation-free entry point (include/rt_b200.h; no counterpart in the reference):
its hit records must equal those of the existing entry point rt_trace_rays,
bit for bit (integer ids; t and N are the same arithmetic), on two calls in a row that share the static device array.
Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rays(oracle_mod, arena, w, h):
    return oracle_mod.camera_rays(arena.camera(), w, h)


def test_intersect_static_matches_trace_rays(rtlib, scene_dir):
    import oracle_binding  # camera_rays only (pure numpy); no oracle library is needed
    ctx = rtlib.Context(0)
    try:
        arena = rtlib.load_ray_file(str(scene_dir / "prims_mixed.ray"))
        ctx.upload(arena)
        p = rtlib.Params.default(depth=5)
        max_n = 4096  # RT_STATIC_MAX_RAYS
        first = _rays(oracle_binding, arena, 64, 48)              # 3072 rays
        second = _rays(oracle_binding, arena, 40, 30)[::-1].copy()  # 1200 other rays: fewer than the first call left behind
        for rays in (first, second):
            assert len(rays) <= max_n
            want, _ = ctx.trace_rays(p, rays, 5, want_rgb=False)
            got = ctx.intersect_static(rays)
            assert (want["obj"] >= 0).sum() > 100
            for f in ("obj", "face", "t", "N"):
                assert np.array_equal(want[f], got[f]), f
        # the full capacity, and an empty call
        full = np.resize(first, (max_n, 6))
        want, _ = ctx.trace_rays(p, full, 5, want_rgb=False)
        got = ctx.intersect_static(full)
        for f in ("obj", "face", "t", "N"):
            assert np.array_equal(want[f], got[f]), f
        assert len(ctx.intersect_static(np.zeros((0, 6)))) == 0
        # beyond the fixed size: refused with a clear message, and the context stays usable
        with pytest.raises(rtlib.RtError) as e:
            ctx.intersect_static(np.resize(first, (max_n + 1, 6)))
        assert e.value.status == 1 and "RT_STATIC_MAX_RAYS" in str(e.value)
        got = ctx.intersect_static(second)
        want, _ = ctx.trace_rays(p, second, 5, want_rgb=False)
        assert np.array_equal(want["obj"], got["obj"]) and np.array_equal(want["t"], got["t"])
    finally:
        ctx.close()


def test_intersect_static_needs_a_scene(rtlib):
    ctx = rtlib.Context(0)
    try:
        with pytest.raises(rtlib.RtError) as e:
            ctx.intersect_static(np.array([[0, 0, 5, 0, 0, -1.0]]))
        assert e.value.status == 4  # RT_ERR_NO_SCENE
    finally:
        ctx.close()
