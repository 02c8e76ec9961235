# ncu --set full of the dominant kernel (first k_wave_shadow_any launch of a C5 frame) on the final build (with the object-space pre-test)
set -x
O=gpurun_out
sum_rep() { python profiles/summarize_ncu.py $O/$1.ncu-rep > $O/$1.txt 2>&1; python profiles/ncu_hot_lines.py $O/$1.ncu-rep >> $O/$1.txt 2>&1; rm -f $O/$1.ncu-rep; }
P="python profiles/part_scaling.py --parts 1 --steps 1 --warmup 0 --workload c5"
$P > $O/r2_final_plain_c5.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:^k_wave_shadow_any\$ -s 0 -c 1 -f -o $O/prof_r2_final_c5_k_wave_shadow_any $P > $O/ncu_k_wave_shadow_any.log 2>&1
sum_rep prof_r2_final_c5_k_wave_shadow_any
