#!/usr/bin/env python3
"""Cost model of the bench workloads: algorithmic bytes and FP64 operations per ray.

The per-ray DEVICE code, compiled for the host with work counters (tests/hostsim, RT_COUNT_WORK),
renders each workload's scene from its camera at a reduced resolution and counts, per tree walk
(one ray = one walk; shadow rays may take more than one collection pass):
    nodes  two-child tree nodes visited   (128 B each: both child boxes + child refs)
    tris   triangle tests                 ( 96 B each: 3 vertices + face normal)
    objs   scene-level candidates tested  (192 B: bbox + inverse transform of the object record)
and the fixed per-ray record traffic of the wavefront (queue entry written + read, hit record,
accumulator update).  These are ALGORITHMIC bytes: what the walk must touch if nothing were cached;
the roofline in bench.py divides them by the measured kernel time.

FP64 operation counts per unit (counted from the source, FMA = 2):
    node: 2 boxes x (6 sub + 6 x (1 mul + 4 fma) + 3 sub) = 2 x 63 = 126 flop, ~40 compares/selects
    tri : plane 5+5+1 div, point 6, three edge tests 3 x (6 sub + 9 + 5) = 60, area 20  ~ 100 flop
    obj : bbox 63 + world->object transform 2 x 24 + normalise 12                        ~ 125 flop

Writes scenes/cost_model.json.  Run here (CPU container), results are committed.
"""
import ctypes
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scenes"))

NODE_B, TRI_B, OBJ_B = 128, 96, 192
NODE_F, TRI_F, OBJ_F = 126, 100, 125
RAY_RECORD_B = {0: 2 * 104 + 16 + 16 + 24, 1: 2 * 80 + 24}  # closest: queue w+r, hit w+r, accum; shadow: queue w+r, accum


def main():
    import bench
    import cs378hgraphics_raytracer_b200 as rt
    import oracle_binding as ob
    d = os.path.join(ROOT, "tests", "hostsim")
    subprocess.check_call(["make", "-s", "-C", d])
    S = ctypes.CDLL(os.path.join(d, "libhostsim.so"))
    S.sim_work.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    S.sim_render.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    out = {}
    keys = sys.argv[1:] or ["c1", "c2", "c3", "c4", "c5"]
    path_json = os.path.join(ROOT, "scenes", "cost_model.json")
    if os.path.exists(path_json):
        out = json.load(open(path_json))
    for key in keys:
        wl = bench.WORKLOADS[key]
        path, faces = bench.scene_files(wl)
        arena = rt.load_ray_file(path, faces)
        cam = arena.camera()
        aspect = cam["aspect"]
        p = rt.Params.default(**{k: v for k, v in wl["params"].items() if k in ("depth", "use_cubemap")})
        w = 192
        h = ob.frame_height(w, aspect)
        buf = np.zeros(2 * 37, np.uint64)
        assert S.sim_work(arena.ptr, ctypes.byref(p), w, h, buf.ctypes.data) == 0
        buf = buf.reshape(2, 37)
        img = np.zeros((h, w, 3), np.uint8)
        st = rt.Stats()
        assert S.sim_render(arena.ptr, ctypes.byref(p), w, h, img.ctypes.data, ctypes.byref(st)) == 0
        n_rays = {0: st.camera_rays + st.reflection_rays + st.refraction_rays, 1: st.shadow_rays}
        kinds = {}
        tot_rays = 0
        tot_bytes = 0.0
        tot_flops = 0.0
        for m, name in ((0, "closest"), (1, "shadow")):
            walks, nodes, tris, objs = [int(x) for x in buf[m, :4]]
            hist = [int(x) for x in buf[m, 4:36]]
            while hist and hist[-1] == 0:
                hist.pop()
            rays = int(n_rays[m])
            kb = nodes * NODE_B + tris * TRI_B + objs * OBJ_B + rays * RAY_RECORD_B[m]
            kf = nodes * NODE_F + tris * TRI_F + objs * OBJ_F
            kinds[name] = dict(rays=rays, walks=walks, nodes_per_ray=nodes / max(rays, 1), tris_per_ray=tris / max(rays, 1),
                               objs_per_ray=objs / max(rays, 1), bytes_per_ray=kb / max(rays, 1), flops_per_ray=kf / max(rays, 1),
                               max_nodes=int(buf[m, 36]), hist_log2_nodes=hist)
            tot_rays += rays
            tot_bytes += kb
            tot_flops += kf
        out[key] = dict(workload=wl["name"], sampled_at="%dx%d, pinhole rays, depth %d" % (w, h, p.depth), rays=tot_rays,
                        bytes_per_ray=tot_bytes / tot_rays, flops_per_ray=tot_flops / tot_rays, kinds=kinds)
        print(key, json.dumps(out[key]))
        sys.stdout.flush()
        with open(path_json, "w") as f:
            json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
