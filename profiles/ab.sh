#!/bin/sh
# A/B of tuning variants on ONE box: sh profiles/ab.sh "<workloads>" <steps> <name> [<name> ...]
#   name = base (the in-tree library) or a variant built by `make -C cs378hgraphics_raytracer_b200/csrc variant NAME=<name> EXTRA_NVFLAGS=...`
# prints: variant workload Mrays/s ms/frame and the per-stage milliseconds
WL="$1"; STEPS="$2"; shift 2
for v in "$@"; do
  for w in $WL; do
    if [ "$v" = base ]; then unset RT_B200_LIB; else export RT_B200_LIB=$PWD/cs378hgraphics_raytracer_b200/variants/librt_b200_$v.so; fi
    python bench.py --workload $w --steps $STEPS --warmup 3 --no-cpu-baseline 2> gpurun_out/ab_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$v', '$w', round(d['value'],1), round(d['ms_per_step'],2), {a:round(b['ms_per_step'],2) for a,b in d['stages'].items()})"
  done
done
