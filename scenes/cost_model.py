#!/usr/bin/env python3
"""Cost model of the bench workloads: SURVEY.md 8(d)'s algorithmic work per ray.

The per-ray DEVICE code, compiled for the host with work counters (tests/hostsim, RT_COUNT_WORK),
renders each workload's scene from its camera at a reduced resolution with the ORDERED, PRUNED walk
(the 8(d) "ordered-traversal lower bound" - not the reference's exhaustive walk) and counts per ray
kind (closest-hit rays = camera + reflection + refraction; shadow rays):
    nodes  two-child tree nodes visited, both tree levels      N_box = walks + 2 * nodes
    tris   triangle tests in entered leaves                      N_tri
    objs   scene-level candidates tested (bbox + transform)      N_obj = N_xform
and applies the fixed formulas of SURVEY.md 8(d):
    bytes/ray = 64 N_box + 128 N_tri + 192 N_obj + 96
    flops/ray = 12 N_box + 150 N_tri + 60 N_obj + 40 N_xform + 120 N_shade_light
(N_shade_light = one per shadow ray, charged to the shadow kind).  These are ALGORITHMIC figures:
what the walk must touch if nothing were cached; bench.py divides them by measured kernel time.

Writes scenes/cost_model.json.  Run here (CPU container); the result is committed.
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

BOX_B, TRI_B, OBJ_B, RAY_B = 64, 128, 192, 96
BOX_F, TRI_F, OBJ_F, XFORM_F, SHADE_F = 12, 150, 60, 40, 120


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
            # per WALKED ray: shadow rays with an exactly zero Phong term are counted but never walked
            # (rt_shade.cuh), bench.py multiplies these figures by the walked rays only
            rays = min(int(n_rays[m]), walks)
            n_box = walks + 2 * nodes
            kb = n_box * BOX_B + tris * TRI_B + objs * OBJ_B + rays * RAY_B
            kf = n_box * BOX_F + tris * TRI_F + objs * (OBJ_F + XFORM_F) + (rays * SHADE_F if m == 1 else 0)
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
