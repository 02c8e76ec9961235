set -x
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r1_v12_bench_c5_short.json 2> gpurun_out/r1_v12_err.log || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1_v12_c5_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
P="python profiles/part_scaling.py --parts 1 --steps 1 --warmup 0"
for k in shadow closest shade; do
  ncu --set full --clock-control none --import-source on -k regex:k_wave_$k\$ -s 5 -c 1 -f -o gpurun_out/prof_v12_c5_$k $P --workload c5 > gpurun_out/ncu_c5_$k.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:k_wave_$k\$ -s 0 -c 1 -f -o gpurun_out/prof_v12_c3_$k $P --workload c3 > gpurun_out/ncu_c3_$k.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
