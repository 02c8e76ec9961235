#!/usr/bin/env python3
"""bench.py - throughput of the trace loop (RayTracer::traceImage) in Mrays/s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl b200|reference]
  (N > 1: launched by torchrun, one rank per GPU; see the contract in the task description)

A step = one full frame of the workload.  rays = camera (incl. AA samples and DOF lens rays) +
reflection + refraction + shadow rays, counted on the device exactly as the instrumented reference
counts them.  `value` is measured with the scene resident in HBM and the frame left
in HBM; `e2e` goes through the host-facing calls every step: rt_upload_scene (host -> device copy of
the whole flattened scene) + render + device -> host copy of the RGB8 frame.

--impl reference times the UNMODIFIED reference (oracle/_ref/libref_oracle.so) on the host cores on
a bounded sample of the same frame.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scenes"))

# BASELINE.json configs.  All scenes are procedural stand-ins (the reference's assets are not in
# its repository); generators and seeds in scenes/gen_scenes.py.
WORKLOADS = {
    "c1": dict(name="C1 hitchcock stand-in 512x512, depth 5, no AA", gen=("hitchcock", 1.0), w=512, params=dict(depth=5)),
    "c2": dict(name="C2 trimesh2 stand-in 1024x1024, depth 5, DOF aperture 0.05 focal 2.5 divs 5", gen=("trimesh2", 1.0), w=1024,
               params=dict(depth=5, dof_on=1, dof_fd=2.5, dof_divs=5, dof_apsz=0.05)),
    "c3": dict(name="C3 dragon stand-in (869,556 faces, kd-tree) 1920x1080, depth 5, shadows", gen=("dragon", 1.0), w=1920,
               params=dict(depth=5)),
    "c4": dict(name="C4 hitchcock stand-in 2048x2048, 4x4 supersampling, cubemap, depth 5", gen=("hitchcock", 1.0), w=2048,
               params=dict(depth=5, aa_mode=1, aa_samples=4, use_cubemap=1), cubemap=512),
    "c4a": dict(name="C4 hitchcock stand-in 2048x2048, adaptive AA 4x4 thresh 0.01, cubemap, depth 5", gen=("hitchcock", 1.0), w=2048,
                params=dict(depth=5, aa_mode=2, aa_samples=4, aa_thresh=0.01, use_cubemap=1), cubemap=512),
    # -O c (adaptive termination) on two configurations: the mode that runs on the fused kernel k_trace_pool
    "c1t": dict(name="C1 hitchcock stand-in 512x512, depth 5, adaptive termination 0.05", gen=("hitchcock", 1.0), w=512, params=dict(depth=5, aterm_thresh=0.05)),
    "c3t": dict(name="C3 dragon stand-in 1920x1080, depth 5, adaptive termination 0.05", gen=("dragon", 1.0), w=1920, params=dict(depth=5, aterm_thresh=0.05)),
    "c5": dict(name="C5 synthetic 1M-primitive mixed scene 7680x4320, depth 8", gen=("mixed", 1.0), w=7680, params=dict(depth=8)),
}
SCENE_CACHE = os.environ.get("RT_SCENE_CACHE", os.path.join(tempfile.gettempdir(), "rt_b200_scenes"))


def scene_files(wl):
    """Generates (once per box) the stand-in .ray text and cube map for a workload."""
    import gen_scenes
    os.makedirs(SCENE_CACHE, exist_ok=True)
    gen, scale = wl["gen"]
    path = os.path.join(SCENE_CACHE, "%s_%g.ray" % (gen, scale))
    if not os.path.exists(path):
        tmp = path + ".tmp%d" % os.getpid()
        with open(tmp, "w") as f:
            f.write(gen_scenes.GENERATORS[gen](scale))
        os.replace(tmp, path)
    faces = None
    if wl.get("cubemap"):
        d = os.path.join(SCENE_CACHE, "cubemap%d" % wl["cubemap"])
        if not os.path.exists(os.path.join(d, "nz.bmp")):
            gen_scenes.gen_cubemap(d, wl["cubemap"])
        faces = [os.path.join(d, n + ".bmp") for n in ("px", "nx", "py", "ny", "pz", "nz")]
    return path, faces


def load_arena(wl):
    """The workload's flattened scene: parsed from the .ray text once per box (9 s for C5), then taken from
    the binary cache next to it (rt_arena_save / rt_arena_load; a stale or damaged cache is refused)."""
    import cs378hgraphics_raytracer_b200 as rt
    path, faces = scene_files(wl)
    cache = path + (".cube%d" % wl["cubemap"] if wl.get("cubemap") else "") + ".arena"
    if os.path.exists(cache):
        try:
            return rt.load_arena(cache)
        except rt.RtError:
            pass
    a = rt.load_ray_file(path, faces)
    try:
        tmp = cache + ".tmp%d" % os.getpid()
        rt.save_arena(tmp, a)
        os.replace(tmp, cache)
    except (rt.RtError, OSError):
        pass
    return a


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md's clocks line).
    NVML is polled from a thread every few milliseconds (the timed region of the small workloads is
    only tens of milliseconds - too short for an nvidia-smi loop); nvidia-smi -lms is the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index
        self.nvml, self.handle, self.stop_flag, self.thread = None, None, False, None
        self.sm, self.reasons, self.max_mhz = [], set(), None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.gpu
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    idx = int(vis.split(",")[self.gpu])
                except (ValueError, IndexError):
                    pass
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _poll(self):
        n = self.nvml
        masks = {"hw_slowdown": getattr(n, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(n, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(n, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(n, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while not self.stop_flag:
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                for name, m in masks.items():
                    if r & m:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml:
            self.stop_flag = True
            self.thread.join(timeout=2)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                pass
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


# ---- the reference on the host cores ------------------------------------------------------------
def reference_sample(orc, params, w, h, budget_s, block=128):
    """Times the unmodified reference's per-pixel code on blocks of the workload's frame, visited
    in a fixed scattered order, until `budget_s` seconds of traceImage-equivalent time are spent
    (or the frame is complete).  Returns (rays, seconds, description)."""
    bx, by = (w + block - 1) // block, (h + block - 1) // block
    order = np.random.default_rng(12345).permutation(bx * by)
    rays, secs, done_px = 0, 0.0, 0
    n_blocks = 0
    for b in order:
        i0, j0 = (b % bx) * block, (b // bx) * block
        i1, j1 = min(i0 + block, w), min(j0 + block, h)
        _, st, s = orc.render_window(params, w, h, i0, i1, j0, j1)
        rays += st.rays
        secs += s
        done_px += (i1 - i0) * (j1 - j0)
        n_blocks += 1
        if secs >= budget_s:
            break
    desc = "%d scattered %dx%d-pixel blocks = %.1f%% of the %dx%d frame" % (n_blocks, block, block, 100.0 * done_px / (w * h), w, h)
    return rays, secs, desc


def run_reference(args, wl, rank):
    """--impl reference: rank 0 alone runs; one step = one bounded sample of the frame."""
    if rank != 0:
        return
    import cs378hgraphics_raytracer_b200 as rt
    import oracle_binding as ob
    if not ob.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_oracle.so missing (reference is built from /root/reference by oracle/Makefile)"}), flush=True)
        return
    arena = load_arena(wl)  # the product's loader only parses; the reference's own load of the dragon takes ~32 min
    orc = ob.Oracle(None, arena=arena)
    params = rt.Params.default(**wl["params"])
    w = wl["w"]
    h = ob.frame_height(w, orc.aspect)
    cores = ob.ref().ref_get_threads()
    budget = max(2.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        reference_sample(orc, params, w, h, budget / 4)
    t_rays, t_secs, desc = 0, 0.0, ""
    for _ in range(args.steps):
        r, s, desc = reference_sample(orc, params, w, h, budget)
        t_rays += r
        t_secs += s
    value = t_rays / t_secs / 1e6
    out = {"metric": "Mrays/s (primary+secondary+shadow)", "value": value, "unit": "Mrays/s", "impl": "reference", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_secs / args.steps, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": wl["name"], "width": w, "height": h, "scene": "procedural stand-in (scenes/gen_scenes.py)"},
           "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": "reference", "sample": desc + " per step"},
           "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    # the dragon frame (targets.c3 of the b200 arm): the whole frame, it takes the host cores about a second
    out["targets"] = {}
    for tkey in [t for t in args.targets.split(",") if t and t != args.workload]:
        twl = WORKLOADS[tkey]
        torc = ob.Oracle(None, arena=load_arena(twl))
        tp = rt.Params.default(**twl["params"])
        tw = twl["w"]
        th = ob.frame_height(tw, torc.aspect)
        full = tkey in ("c1", "c2", "c3")
        r, s, desc = reference_sample(torc, tp, tw, th, 1e9 if full else budget)
        out["targets"][tkey] = {"value": r / s / 1e6, "unit": "Mrays/s", "ms_per_step": 1e3 * s, "cores": cores, "sample": desc,
                                "config": {"workload": twl["name"], "width": tw, "height": th},
                                "e2e": {"value": r / s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


# ---- the B200 arm --------------------------------------------------------------------------------
def load_cost_model(key):
    """Algorithmic bytes / flops per ray of a workload (SURVEY.md 8d cost model), computed once by
    scenes/cost_model.py with the instrumented host-sim and stored next to the scenes."""
    p = os.path.join(ROOT, "scenes", "cost_model.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(key)
    return None


def load_dram_traffic(key, kind):
    """DRAM bytes of one launch of the stage kernel from the committed `ncu --set full` capture."""
    p = os.path.join(ROOT, "profiles", "dram_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return (json.load(f).get(key) or {}).get(kind)
    return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f).get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measure(S, key, wl, steps, warmup, cpu_seconds, cpu_full_frame):
    """One workload on the session's GPUs: the bench line's fields for it (rank 0 returns the dict, the others None).
    S = session (ctx, rank, world, device, dist, modules)."""
    import torch
    rt, par, ob, ctx, rank, world, device, dist = S["rt"], S["par"], S["ob"], S["ctx"], S["rank"], S["world"], S["device"], S["dist"]
    params = rt.Params.default(**wl["params"])

    # scene: rank 0 parses + flattens, NCCL broadcast replicates it (north star: scene replicated)
    arena_bytes = b""
    if rank == 0:
        a = load_arena(wl)
        arena_bytes = a.bytes()
        a.free()
    arena_bytes = par.broadcast_arena(arena_bytes, rank, world, device)
    # the arena lives in page-locked host memory (the e2e step copies it to the device every time)
    pinned = torch.empty(len(arena_bytes), dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[...] = np.frombuffer(arena_bytes, dtype=np.uint8)

    def upload():
        ctx._check(rt.lib().rt_upload_scene(ctx._h, ctypes.c_void_p(pinned.data_ptr()), pinned.numel()))

    upload()
    hdr = np.frombuffer(arena_bytes[:32 + 17 * 24], dtype=np.uint64)
    cam = np.frombuffer(arena_bytes[int(hdr[4 + 3 * 1 + 1]):int(hdr[4 + 3 * 1 + 1]) + 14 * 8], dtype=np.float64)
    w = wl["w"]
    h = ob.frame_height(w, float(cam[12]))
    h2d_bytes = len(arena_bytes)
    flush = S["flush"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        flush.fill_(1)
        torch.cuda.synchronize()
        ptr, st = ctx.render_device(params, w, h, rank, world)
        if world > 1:
            par.gather_frame_device(ctx, ptr, w, h, rank, world, device)
            torch.cuda.synchronize()
        return st

    host_pinned = torch.empty((h, w, 3), dtype=torch.uint8, pin_memory=True)  # the caller's frame buffer, page-locked
    host_out = host_pinned.numpy()

    def step_e2e():
        upload()
        if world == 1:
            _, st = ctx.render(params, w, h, out=host_out)
        else:
            ptr, st = ctx.render_device(params, w, h, rank, world)
            full = par.gather_frame_device(ctx, ptr, w, h, rank, world, device)
            if rank == 0:
                host_pinned.copy_(full, non_blocking=True)  # page-locked destination: one DMA, no staging copy
            torch.cuda.synchronize()
        return st

    for _ in range(warmup):
        step_resident()
    sampler = ClockSampler(S["local_rank"])
    if rank == 0:
        sampler.start()
    barrier()
    t0 = time.perf_counter()
    rays = launches = culled = 0
    dev_ms = 0.0
    for _ in range(steps):
        st = step_resident()
        rays += st.rays
        culled += st.culled_shadow_rays
        launches += st.launches
        dev_ms += st.device_ms
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop() if rank == 0 else None
    # the L2 flush is not part of the step: measure it alone and take it out
    torch.cuda.synchronize()
    f0 = time.perf_counter()
    for _ in range(steps):
        flush.fill_(1)
        torch.cuda.synchronize()
    flush_s = time.perf_counter() - f0
    elapsed = max(1e-9, (t1 - t0) - flush_s)

    step_e2e()
    barrier()
    e0 = time.perf_counter()
    e_rays = 0
    for _ in range(steps):
        e_rays += step_e2e().rays
    barrier()
    e1 = time.perf_counter()
    # per-stage device times: the same frames once more with CUDA events around every launch
    # (rt_set_profiling); kept out of the timed region above because the extra events cost a few
    # microseconds per launch and the stages are serialised for clean times
    ctx.set_profiling(True)
    stage_ms = np.zeros(6)
    stage_n = np.zeros(6)
    kind_rays = np.zeros(2)
    for _ in range(steps):
        st = step_resident()
        stage_ms += np.array(list(st.stage_ms))
        stage_n += np.array(list(st.stage_launches))
        # rays the two walk kernels really walked: shadow rays dropped by k_wave_shade (exactly zero Phong term) are not among them
        kind_rays += np.array([st.camera_rays + st.reflection_rays + st.refraction_rays, st.shadow_rays - st.culled_shadow_rays], dtype=np.float64)
    ctx.set_profiling(False)

    stats = torch.tensor([float(rays), float(e_rays), float(launches), elapsed, e1 - e0, dev_ms, float(culled)], dtype=torch.float64, device=device)
    if world > 1:
        sums, maxs, mins = stats.clone(), stats.clone(), stats.clone()
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
        dist.all_reduce(maxs, op=dist.ReduceOp.MAX)
        dist.all_reduce(mins, op=dist.ReduceOp.MIN)
    else:
        sums = maxs = mins = stats
    if rank != 0:
        return None
    tot_rays, tot_e_rays, tot_launches, tot_culled = float(sums[0]), float(sums[1]), int(sums[2]), float(sums[6])
    t_max, e_max, dev_ms_max, dev_ms_min = float(maxs[3]), float(maxs[4]), float(maxs[5]), float(mins[5])
    value = tot_rays / t_max / 1e6
    out = {"metric": "Mrays/s (primary+secondary+shadow)", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": steps,
           "warmup": warmup, "ms_per_step": 1e3 * t_max / steps, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           # `config` is the same object on both arms (--impl reference); what depends on the run is in config_notes
           "config": {"workload": wl["name"], "width": w, "height": h, "scene": "procedural stand-in (scenes/gen_scenes.py)"},
           "config_notes": {"rays_per_frame": tot_rays / steps, "scene": "the reference's own asset is not in its repository (SURVEY.md 0.3)",
                      "partition": "interleaved 32x32 tiles, tile t -> rank t %% %d; scene replicated" % world,
                      "l2": "256 MiB L2 flush between timed frames (its own time subtracted)",
                      "rays": "counted as the instrumented reference counts them; %d shadow rays per frame are counted but not walked (Phong term exactly 0: their contribution is 0 whatever they hit) - value_walked leaves them out" % (tot_culled // max(1, steps))},
           # the same throughput over the rays the GPU really walks (what the roofline below is about)
           "value_walked": (tot_rays - tot_culled) / t_max / 1e6,
           # CUDA events on the launching stream around each frame's device work; max and min over ranks tell tile imbalance
           # (max >> min) from round tails (both far above the N = 1 time / N); the gather is outside them
           "device_ms_per_step": dev_ms_max / steps, "device_ms_per_step_min_rank": dev_ms_min / steps,
           "clocks": clocks, "gpu_launches": tot_launches,
           "e2e": {"value": tot_e_rays / e_max / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": w * h * 3,
                   "what": "rt_upload_scene (whole scene arena, pinned host -> device, layouts rebuilt on the device) + render + RGB8 frame device -> host, every step"}}
    # roofline of the dominant kernel (rank 0's stage times)
    cm = load_cost_model(key)
    peak, peak_src = measured_peaks()
    fp64_peak, fp32_peak = ctx.measure_fma_peak()
    names = rt.Stats.STAGES
    out["stages"] = {names[k]: {"ms_per_step": stage_ms[k] / steps, "launches_per_step": stage_n[k] / steps} for k in range(6) if stage_n[k] > 0}
    if cm and stage_ms.sum() > 0:
        closest_ms, shadow_ms = stage_ms[1] + stage_ms[2], stage_ms[4]
        kind, kernel, k_ms, k_n = ("shadow", "k_wave_shadow_any (+ k_wave_shadow_list)", shadow_ms, stage_n[4]) if shadow_ms >= closest_ms else \
                                  ("closest", "k_wave_closest (+ k_wave_closest_heavy)", closest_ms, stage_n[1])
        ck = cm["kinds"][kind]
        n_rays = kind_rays[0 if kind == "closest" else 1]
        alg_bytes = ck["bytes_per_ray"] * n_rays
        ach = alg_bytes / (k_ms * 1e-3) / 1e9
        # bytes of the records the kernel really reads for the same visits: 64 B float node (both children), 32 B float item
        # box, 96 B triangle, 248 B object (inverse, normal matrix, box) + 96 B ray in / result out
        lay_bytes = (64.0 * ck["nodes_per_ray"] + 32.0 * ck["objs_per_ray"] + 96.0 * ck["tris_per_ray"] + 248.0 * min(ck["objs_per_ray"], 3.0) + 96.0) * n_rays
        # DRAM bytes and counters of this kernel from the committed ncu capture of one launch
        tr = load_dram_traffic(key, kind)
        traffic, traffic_src, cnt = None, None, {}
        if isinstance(tr, dict) and tr.get("launch_ms"):
            traffic = tr["bytes"] * (k_ms / max(k_n, 1)) / tr["launch_ms"]
            traffic_src = "%s; scaled from its %.2f ms to this run's %.2f ms per launch" % (tr.get("source", ""), tr["launch_ms"], k_ms / max(k_n, 1))
            cnt = tr.get("counters", {})
        fp64_ach = ck["flops_per_ray"] * n_rays / (k_ms * 1e-3) / 1e12
        out["roofline"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                           "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel": kernel,
                           "kernel_ms_per_launch": k_ms / max(k_n, 1), "launches_per_step": k_n / steps,
                           "algorithmic_bytes_per_launch": alg_bytes / max(k_n, 1), "bytes_per_ray": ck["bytes_per_ray"],
                           "numerator": "SURVEY 8(d) formula (64 B per box, 128 B per triangle, 192 B per object, 96 B per ray) over the DEVICE walk's own visit "
                                        "counts (host build of the device headers with counters, scenes/cost_model.py) - not an oracle-derived lower bound",
                           "kernel_layout_bytes_per_launch": lay_bytes / max(k_n, 1), "kernel_layout_frac": lay_bytes / (k_ms * 1e-3) / 1e9 / peak,
                           # what the counters say (ncu capture of one launch of this kernel, profiles/): none of the three roofs is near
                           "dram_frac": (traffic / (k_ms / max(k_n, 1) * 1e-3) / 1e9 / peak) if traffic else None,
                           "fp64_frac": fp64_ach / fp64_peak if fp64_peak else None,
                           "lanes_per_inst": cnt.get("lanes_per_inst"), "issue_active_pct": cnt.get("issue_active_pct"),
                           "limiter": "latency and divergence: the scene is cache resident, DRAM and FP64 are idle; the walk is a chain of dependent L2 accesses at a fraction of the 32 lanes (DESIGN.md section 5)",
                           "fp64": {"flops_per_ray": ck["flops_per_ray"], "achieved_tflops": fp64_ach, "peak_tflops": fp64_peak,
                                    "peak_source": "measured: rt_measure_fma_peak (8 FMA chains per thread)", "fp32_peak_tflops": fp32_peak}}
    else:
        out["roofline"] = None
    # the reference on this box's host cores (N = 1 only): a bounded sample, or the whole frame where that takes seconds
    if world == 1 and ob.available() and cpu_seconds > 0:
        arena = load_arena(wl)
        orc = ob.Oracle(None, arena=arena)
        r, s, desc = reference_sample(orc, params, w, h, 1e9 if cpu_full_frame else cpu_seconds)
        out["cpu_baseline"] = {"value": r / s / 1e6, "unit": "Mrays/s", "cores": ob.ref().ref_get_threads(), "kind": "reference", "sample": desc}
    else:
        out["cpu_baseline"] = None
    return out


def run_b200(args, key, wl, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import cs378hgraphics_raytracer_b200 as rt
    from cs378hgraphics_raytracer_b200 import parallel as par
    import oracle_binding as ob

    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    S = {"rt": rt, "par": par, "ob": ob, "ctx": rt.Context(local_rank), "rank": rank, "world": world, "device": device, "dist": dist,
         "local_rank": local_rank, "flush": torch.empty(256 << 20, dtype=torch.uint8, device=device)}  # flush buffer > 126 MB L2
    cpu_s = 0.0 if args.no_cpu_baseline else args.cpu_seconds
    out = measure(S, key, wl, args.steps, args.warmup, cpu_s, False)
    # the scene the north star's target names (the dragon, C3), measured in the same run and - under torchrun - at the same N,
    # so that the driver's BENCH / SCALE records carry it next to the headline workload
    targets = {}
    for tkey in [t for t in args.targets.split(",") if t and t != key]:
        twl = WORKLOADS[tkey]
        t = measure(S, tkey, twl, max(5, args.steps), max(3, args.warmup), cpu_s, tkey in ("c1", "c2", "c3"))
        if t is not None:
            targets[tkey] = {k: t[k] for k in ("value", "value_walked", "unit", "ms_per_step", "device_ms_per_step", "device_ms_per_step_min_rank", "steps", "warmup",
                                               "gpu_launches", "e2e", "roofline", "cpu_baseline", "stages", "config", "config_notes")}
    if rank == 0:
        out["targets"] = targets
        print(json.dumps(out), flush=True)
    S["ctx"].close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    # default = the configuration BASELINE.json's metric ("Mrays/s at 1/2/4/8 B200") is quoted on: configs[4], the one
    # "tile-partitioned across 1/2/4/8 GPUs"; it fits one GPU.  c3 (the dragon) is the scene the target ratio names.
    ap.add_argument("--workload", default="c5", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="host-core time spent on the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--targets", default="c3", help="comma list of further workloads measured in the same run and reported under `targets` (the dragon scene the north star's ratio names); '' = none")
    args = ap.parse_args()
    # stdout carries ONE JSON line and nothing else: libraries that write to file descriptor 1 (NCCL prints
    # "NCCL version ..." there when NCCL_DEBUG is set) are sent to stderr, the line goes to the saved descriptor
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl, rank)
    else:
        run_b200(args, args.workload, wl, rank, world, local_rank)


if __name__ == "__main__":
    main()
