# final build of round 1: launch list of the default bench command, full captures of the two walk kernels on C5
set -x
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r1_v14_bench_c5_short.json 2> gpurun_out/r1_v14_err.log || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_v14_c5_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
P="python profiles/part_scaling.py --parts 1 --steps 1 --warmup 0 --workload c5"
for k in shadow closest; do
  ncu --set full --clock-control none --import-source on -k regex:k_wave_$k\$ -s 1 -c 1 -f -o gpurun_out/prof_v14_c5_$k $P > gpurun_out/ncu_c5_$k.log 2>&1
done
ls -la gpurun_out/prof_v14*.ncu-rep
