# The build with the object-space float pre-test of scene-level candidates (commit e1ac7c2): GPU tests, same-box A/B against the build
# before it (variant "head" = commit df95c09), then the bench lines of the final build.
set -x
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $O/r2_pretest_tests.txt
sh profiles/ab.sh "c5 c3" 3 base head > $O/r2_pretest_ab.txt 2>&1
python bench.py > $O/r2_final_bench_default.json 2> $O/r2_final_bench_default.err
python bench.py --workload c4 --targets c4a,c1,c2,c3 --cpu-seconds 5 > $O/r2_final_bench_c4_and_targets.json 2> $O/r2_final_bench_c4.err
