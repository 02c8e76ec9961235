#!/usr/bin/env python3
"""Lane-utilisation model of the first round of a frame (CPU only, tests/hostsim with work counters).

For every pixel of a depth-0 frame: tree nodes visited by the camera ray's walk and by each of its
shadow walks.  A warp runs as long as its slowest lane, so for an assignment of rays to warps
    efficiency = sum(nodes) / (32 * sum over warps of max(nodes in warp))
which is what ncu reports as "active threads per warp-instruction" / 32 for the walk kernels.
Compares assignments: row order (32x1 pixels per warp), 8x4 pixel tiles, and an ideal sort by length.
usage: python profiles/simt_model.py c3 [width]
"""
import ctypes
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def eff(a):
    a = np.asarray(a, np.float64)
    n = len(a) // 32 * 32
    g = a[:n].reshape(-1, 32)
    return a[:n].sum() / max(1.0, 32.0 * g.max(axis=1).sum())


def capped(a, budget):
    """walks longer than the budget leave the warp after `budget` nodes (cooperative re-walk)"""
    return np.minimum(a, budget)


def tiles(img, tw, th):
    h, w = img.shape
    hh, ww = h // th * th, w // tw * tw
    return img[:hh, :ww].reshape(hh // th, th, ww // tw, tw).transpose(0, 2, 1, 3).reshape(-1)


def main():
    import bench
    import cs378hgraphics_raytracer_b200 as rt
    import oracle_binding as ob
    key = sys.argv[1] if len(sys.argv) > 1 else "c3"
    w = int(sys.argv[2]) if len(sys.argv) > 2 else 480
    d = os.path.join(ROOT, "tests", "hostsim")
    subprocess.check_call(["make", "-s", "-C", d])
    S = ctypes.CDLL(os.path.join(d, "libhostsim.so"))
    S.sim_walk_lengths.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    wl = bench.WORKLOADS[key]
    path, faces = bench.scene_files(wl)
    arena = rt.load_ray_file(path, faces)
    h = ob.frame_height(w, arena.camera()["aspect"])
    p = rt.Params.default(depth=0)
    slots = 1 + 4
    out = np.zeros((h, w, slots), np.uint32)
    assert S.sim_walk_lengths(arena.ptr, ctypes.byref(p), w, h, slots, out.ctypes.data) == 0
    print("%s at %dx%d (depth 0)" % (key, w, h))
    for s, name in [(0, "camera walk")] + [(k, "shadow walk %d" % (k - 1)) for k in range(1, slots)]:
        a = out[:, :, s]
        if a.sum() == 0:
            continue
        live = a > 0
        print("%-14s rays %8d  mean nodes %7.1f  p50 %5d p90 %5d p99 %6d max %7d" %
              (name, live.sum(), a[live].mean() if live.any() else 0, *np.percentile(a[live], [50, 90, 99]).astype(int), a.max()))
        for budget in (1 << 30, 1024, 256, 96):
            b = capped(a, budget)
            print("   budget %-10s row order %.3f   8x4 tiles %.3f   4x8 tiles %.3f   sorted %.3f" %
                  ("none" if budget == 1 << 30 else budget, eff(b.reshape(-1)), eff(tiles(b, 8, 4)), eff(tiles(b, 4, 8)), eff(np.sort(b.reshape(-1)))))
    # shadow queue as the shade stage builds it: per warp of hits, light by light (only lit pixels)
    for order_name, order in (("row", lambda x: x.reshape(-1)), ("8x4", lambda x: tiles(x, 8, 4))):
        hit = order(out[:, :, 0] * 0 + (out[:, :, 1:].sum(axis=2) > 0))
        qs = []
        cols = [order(out[:, :, k]) for k in range(1, slots) if out[:, :, k].sum() > 0]
        n = len(hit) // 32 * 32
        for g in range(0, n, 32):
            m = hit[g:g + 32] > 0
            for c in cols:
                qs.append(c[g:g + 32][m])
        q = np.concatenate(qs) if qs else np.zeros(0)
        print("shadow queue (%s order of hits, light by light): %d rays, efficiency %.3f, budget 256 %.3f, sorted %.3f" %
              (order_name, len(q), eff(q), eff(capped(q, 256)), eff(np.sort(q))))


if __name__ == "__main__":
    main()
