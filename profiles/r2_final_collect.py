#!/usr/bin/env python3
"""Collects what `r2_final_capture.sh` left under gpurun_out/ into profiles/:
  * the bench lines (r2_final_bench_*.json) and the launch list (r2_final_c5_launches.csv), copied;
  * the per-kernel ncu summaries prof_r2_final_<wl>_<kernel>.txt -> r2_final_<wl>_<kernel>.txt;
  * r2_final_c5_launch_shares.txt: share of each kernel in the launch list, next to the stage times of the same command
    run without ncu (r2_final_bench_c5_short.json);
  * dram_traffic.json: DRAM bytes, duration and headline counters of the captured launches of the two walk kernels,
    which bench.py scales to its own average launch (roofline.traffic).
Run here after the gpurun call:  python profiles/r2_final_collect.py"""
import csv
import glob
import json
import os
import re
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "gpurun_out")
DST = os.path.join(ROOT, "profiles")


def metrics(path):
    """metric name -> (unit, value) of the first kernel section of a summarize_ncu.py text"""
    out = {}
    for line in open(path):
        m = re.match(r"^([a-z_0-9.]+)\s+(\S+)?\s+([-0-9.e+]+)\s*$", line.rstrip())
        if m:
            out.setdefault(m.group(1), (m.group(2) or "", float(m.group(3))))
    return out


def to_bytes(unit, v):
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}[unit]


def to_ms(unit, v):
    return v * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}[unit]


def main():
    for f in glob.glob(os.path.join(SRC, "r2_final_bench_*.json")) + glob.glob(os.path.join(SRC, "r2_final_c5_launches.csv")):
        if os.path.getsize(f) > 0:
            shutil.copy(f, DST)
    for f in glob.glob(os.path.join(SRC, "prof_r2_final_*.txt")):
        shutil.copy(f, os.path.join(DST, os.path.basename(f)[len("prof_"):]))

    # launch shares
    rows = []
    with open(os.path.join(DST, "r2_final_c5_launches.csv")) as fh:
        lines = [l for l in fh if l.startswith('"')]
    rd = csv.DictReader(lines)
    per = {}
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"^(void )?(rtd::)?", "", r["Kernel Name"]).split("(")[0].split("<")[0]
        v = float(r["Metric Value"].replace(",", ""))
        ms = to_ms({"nsecond": "ns", "usecond": "us", "msecond": "ms", "second": "s", "ns": "ns", "us": "us", "ms": "ms"}[r["Metric Unit"]], v)
        n, t = per.get(name, (0, 0.0))
        per[name] = (n + 1, t + ms)
    total = sum(t for _, t in per.values())
    short = json.loads(open(os.path.join(DST, "r2_final_bench_c5_short.json")).read())
    st = {k: v["ms_per_step"] for k, v in short["stages"].items()}
    st_total = sum(st.values())
    with open(os.path.join(DST, "r2_final_c5_launch_shares.txt"), "w") as out:
        out.write('# launch list of `python bench.py --steps 2 --warmup 1 --no-cpu-baseline --targets ""` under ncu (r2_final_c5_launches.csv): share of the summed kernel time\n')
        out.write("# kernel, launches, ms, share\n")
        for name, (n, t) in sorted(per.items(), key=lambda kv: -kv[1][1]):
            out.write("%-36s %4d %10.2f ms  %5.1f %%\n" % (name[:36], n, t, 100 * t / total))
        out.write("# the same command without ncu (rt_set_profiling stage times, ms per frame): " +
                  ", ".join("%s %.1f (%.1f %%)" % (k, v, 100 * v / st_total) for k, v in st.items()) + "\n")
        groups = {"shadow": ("k_wave_shadow",), "closest": ("k_wave_closest",), "shade": ("k_wave_shade",)}
        out.write("# shares, ncu launch list / stage times: " + ", ".join(
            "%s %.1f %% / %.1f %%" % (g, 100 * sum(t for nme, (_, t) in per.items() if nme.startswith(pre) and (g != "closest" or "heavy" not in nme)) / total,
                                      100 * st.get(g, 0.0) / st_total) for g, (pre,) in groups.items()) + "\n")

    # dram_traffic.json
    d = {"_what": "dram__bytes_read.sum + dram__bytes_write.sum of ONE launch from an `ncu --set full` capture of the final round-2 build (cold cache, "
                  "serialised; profiles/r2_final_capture.sh), per workload and stage kernel, with that launch's duration under ncu and its headline "
                  "counters; launches of a frame differ in size (the rays of a round), so bench.py reports roofline.traffic for its AVERAGE launch: "
                  "bytes * (its measured ms per launch / launch_ms)"}
    for wl in ("c5", "c3"):
        d[wl] = {}
        for kind, kernel in (("shadow", "k_wave_shadow_any"), ("closest", "k_wave_closest")):
            path = os.path.join(DST, "r2_final_%s_%s.txt" % (wl, kernel))
            m = metrics(path)
            rd_b = to_bytes(*m["dram__bytes_read.sum"])
            wr_b = to_bytes(*m["dram__bytes_write.sum"])
            d[wl][kind] = {
                "bytes": rd_b + wr_b, "launch_ms": to_ms(*m["gpu__time_duration.sum"]),
                "source": "profiles/r2_final_%s_%s.txt (first launch of the frame, %d blocks, %d registers: %.2f GB read + %.2f GB written)" % (
                    wl, kernel, int(m["launch__grid_size"][1]), int(m["launch__registers_per_thread"][1]), rd_b / 1e9, wr_b / 1e9),
                "counters": {"lanes_per_inst": m["smsp__thread_inst_executed_per_inst_executed.ratio"][1],
                             "issue_active_pct": m["smsp__issue_active.avg.pct_of_peak_sustained_active"][1],
                             "l1_hit_pct": m["l1tex__t_sector_hit_rate.pct"][1], "l2_hit_pct": m["lts__t_sector_hit_rate.pct"][1],
                             "warps_active_pct": m["sm__warps_active.avg.pct_of_peak_sustained_active"][1]}}
    with open(os.path.join(DST, "dram_traffic.json"), "w") as out:
        json.dump(d, out, indent=1)
        out.write("\n")
    print(open(os.path.join(DST, "r2_final_c5_launch_shares.txt")).read())
    print(json.dumps({k: {kk: (vv["bytes"], vv["launch_ms"], vv["counters"]["lanes_per_inst"]) for kk, vv in v.items()} for k, v in d.items() if k != "_what"}))


if __name__ == "__main__":
    main()
