#!/usr/bin/env python3
"""What one rank of an N-way tile partition costs, measured on ONE GPU:
    python profiles/part_scaling.py [--workload c3] [--steps 10]
renders partition 0 of N (N = 1, 2, 4, 8) of the workload's frame through rt_render_device and prints
device ms, wall ms, rounds, launches and per-stage times.  ideal = the N = 1 time / N; the gap is the
fixed cost per round (launches, the host's round planning) plus the tail of the slowest rays."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scenes"))
import bench  # noqa: E402
import cs378hgraphics_raytracer_b200 as rt  # noqa: E402
import oracle_binding as ob  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--parts", default="1,2,4,8")
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    wl = bench.WORKLOADS[args.workload]
    arena = rt.load_ray_file(*bench.scene_files(wl))
    ctx = rt.Context(0)
    ctx.upload(arena)
    params = rt.Params.default(**wl["params"])
    w = wl["w"]
    ab = arena.bytes()
    hdr = np.frombuffer(ab[:32 + 17 * 24], dtype=np.uint64)  # arena header: section table (include/rt_arena.h), as in bench.py
    cam = np.frombuffer(ab[int(hdr[4 + 3 * 1 + 1]):int(hdr[4 + 3 * 1 + 1]) + 14 * 8], dtype=np.float64)
    h = ob.frame_height(w, float(cam[12]))
    base = None
    for n in [int(x) for x in args.parts.split(",")]:
        for _ in range(args.warmup):
            ctx.render_device(params, w, h, 0, n)
        dev, wall, rays = [], [], 0
        for _ in range(args.steps):
            t0 = time.perf_counter()
            _, st = ctx.render_device(params, w, h, 0, n)
            wall.append(1e3 * (time.perf_counter() - t0))
            dev.append(st.device_ms)
            rays = st.rays
        ctx.set_profiling(True)
        _, st = ctx.render_device(params, w, h, 0, n)
        ctx.set_profiling(False)
        stages = {rt.Stats.STAGES[k]: round(st.stage_ms[k], 3) for k in range(6) if st.stage_launches[k]}
        d, wl_ms = float(np.median(dev)), float(np.median(wall))
        if base is None:
            base = wl_ms
        print(json.dumps({"workload": args.workload, "n_parts": n, "rays": rays, "device_ms": round(d, 3), "wall_ms": round(wl_ms, 3),
                          "ideal_ms": round(base / n, 3), "efficiency": round(base / n / wl_ms, 3), "rounds": st.rounds,
                          "launches": st.launches, "stage_ms_serialised": stages}))
    ctx.close()


if __name__ == "__main__":
    main()
