# After the accumulator read-back was made cheap (acc_merge: clz + constructed power of two; k_resolve_grid 1.32 -> see below), the bench
# lines of the final build once more, and the capture of k_resolve_grid.  All other kernels are byte-identical to r2_final_capture.sh's build.
set -x
O=gpurun_out
sum_rep() { python profiles/summarize_ncu.py $O/$1.ncu-rep > $O/$1.txt 2>&1; python profiles/ncu_hot_lines.py $O/$1.ncu-rep >> $O/$1.txt 2>&1; rm -f $O/$1.ncu-rep; }
python bench.py > $O/r2_final_bench_default.json 2> $O/r2_final_bench_default.err || exit 1
python bench.py --workload c4 --targets c4a,c1,c2,c3 --cpu-seconds 8 > $O/r2_final_bench_c4_and_targets.json 2> $O/r2_final_bench_c4.err
P="python profiles/part_scaling.py --parts 1 --steps 1 --warmup 0 --workload c5"
ncu --set full --clock-control none --import-source on -k regex:^k_resolve_grid\$ -s 0 -c 1 -f -o $O/prof_r2_final_c5_k_resolve_grid $P > $O/ncu_k_resolve_grid.log 2>&1
sum_rep prof_r2_final_c5_k_resolve_grid
