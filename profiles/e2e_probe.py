#!/usr/bin/env python3
"""Where an end-to-end step spends its time: rt_upload_scene / render into a page-locked frame buffer.
    python profiles/e2e_probe.py [workload]"""
import ctypes
import os
import sys
import time

sys.path.insert(0, os.getcwd()); sys.path.insert(0, "tests"); sys.path.insert(0, "scenes")
import numpy as np
import torch
import bench
import cs378hgraphics_raytracer_b200 as rt
import oracle_binding as ob

wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "c5"]
a = bench.load_arena(wl)
ab = a.bytes()
pinned = torch.empty(len(ab), dtype=torch.uint8, pin_memory=True)
pinned.numpy()[...] = np.frombuffer(ab, dtype=np.uint8)
ctx = rt.Context(0)
p = rt.Params.default(**wl["params"])
w = wl["w"]
h = ob.frame_height(w, a.camera()["aspect"])
out = torch.empty((h, w, 3), dtype=torch.uint8, pin_memory=True).numpy()


def up():
    ctx._check(rt.lib().rt_upload_scene(ctx._h, ctypes.c_void_p(pinned.data_ptr()), pinned.numel()))


up()
ctx.render(p, w, h, out=out)
tu = tr = td = 0.0
n = 4
for _ in range(n):
    t0 = time.perf_counter(); up(); t1 = time.perf_counter()
    _, st = ctx.render(p, w, h, out=out); t2 = time.perf_counter()
    tu += t1 - t0; tr += t2 - t1; td += st.device_ms
ctx.set_profiling(True)
up()
_, st = ctx.render(p, w, h, out=out)
print("after upload :", st.rounds, "rounds", st.launches, "launches", {rt.Stats.STAGES[k]: round(st.stage_ms[k], 1) for k in range(6) if st.stage_launches[k]}, "cooperative rays", st.cooperative_rays)
_, st = ctx.render(p, w, h, out=out)
print("resident     :", st.rounds, "rounds", st.launches, "launches", {rt.Stats.STAGES[k]: round(st.stage_ms[k], 1) for k in range(6) if st.stage_launches[k]}, "cooperative rays", st.cooperative_rays)
ctx.set_profiling(False)
print("%s: upload %.1f ms, render call %.1f ms (device %.1f ms) per step" % (sys.argv[1] if len(sys.argv) > 1 else "c5", 1e3 * tu / n, 1e3 * tr / n, td / n))
