# Final build of round 2 on one B200: tests, bench lines of every workload, the reference arm, the launch list of the default bench
# command, and one `ncu --set full` capture per kernel of the trace loop.   sh profiles/r2_final_capture.sh   (under gpurun)
set -x
O=gpurun_out
# .ncu-rep files are summarised here and deleted: gpurun brings back at most 64 MiB
sum_rep() { python profiles/summarize_ncu.py $O/$1.ncu-rep > $O/$1.txt 2>&1; python profiles/ncu_hot_lines.py $O/$1.ncu-rep >> $O/$1.txt 2>&1; rm -f $O/$1.ncu-rep; }
python bench.py > $O/r2_final_bench_default.json 2> $O/r2_final_bench_default.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_final_bench_ref.json 2> $O/r2_final_bench_ref.err
for w in c1 c2 c3 c4 c4a; do python bench.py --workload $w --targets "" > $O/r2_final_bench_$w.json 2> $O/r2_final_bench_$w.err; done
# launch list: the same short command first without ncu
S="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --targets ''"
sh -c "$S" > $O/r2_final_bench_c5_short.json 2> $O/r2_final_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2_final_c5_launches.csv sh -c "$S" > $O/ncu_list.log 2>&1
# one full capture per kernel (the first launch of each in a C5 frame; the adaptive-AA kernels on C4a)
P="python profiles/part_scaling.py --parts 1 --steps 1 --warmup 0 --workload c5"
$P > $O/r2_final_plain_c5.log 2>&1 || exit 1
for k in k_wave_shadow_any k_wave_shadow_list k_wave_closest k_wave_shade k_wave_closest_retry k_wave_closest_heavy k_wave_gen k_resolve_grid k_shadow_keys k_closest_keys; do
  ncu --set full --clock-control none --import-source on -k regex:^$k\$ -s 0 -c 1 -f -o $O/prof_r2_final_c5_$k $P > $O/ncu_$k.log 2>&1
  sum_rep prof_r2_final_c5_$k
done
# the one-kernel walk (what the undecided rounds use): forced on
RT_B200_ANYHIT=0 ncu --set full --clock-control none --import-source on -k regex:^k_wave_shadow\$ -s 0 -c 1 -f -o $O/prof_r2_final_c5_k_wave_shadow $P > $O/ncu_k_wave_shadow.log 2>&1
sum_rep prof_r2_final_c5_k_wave_shadow
P3="python profiles/part_scaling.py --parts 1 --steps 1 --warmup 0 --workload c3"
$P3 > $O/r2_final_plain_c3.log 2>&1 && for k in k_wave_shadow_any k_wave_closest; do
  ncu --set full --clock-control none --import-source on -k regex:^$k\$ -s 0 -c 1 -f -o $O/prof_r2_final_c3_$k $P3 > $O/ncu_c3_$k.log 2>&1
  sum_rep prof_r2_final_c3_$k
done
PA="python profiles/part_scaling.py --parts 1 --steps 1 --warmup 0 --workload c4a"
$PA > $O/r2_final_plain_c4a.log 2>&1 && for k in k_adapt_resolve k_adapt_samples; do
  ncu --set full --clock-control none --import-source on -k regex:^$k\$ -s 0 -c 1 -f -o $O/prof_r2_final_c4a_$k $PA > $O/ncu_c4a_$k.log 2>&1
  sum_rep prof_r2_final_c4a_$k
done
ls $O/prof_r2_final_*.txt | wc -l; du -sh $O
