#!/bin/bash
# GPU call B of round 2: overflow-path diagnostics, cost-table collapse A/B, checked build, GPU tests, ncu captures (C3 k_intersect, C2 k_extend / k_shade)
mkdir -p gpurun_out
V=vermilion_b200/variants
for lib in tinystack tiny4 sh2; do for w in 8 2; do
  echo "== $lib bvh$w" >> gpurun_out/b_diag.txt
  VMN_LIB=$V/libvermilion_b200_$lib.so VMN_DEBUG_NO_DEPTH_CHECK=1 timeout 60 python tools/diag_tinystack.py $w 64 >> gpurun_out/b_diag.txt 2>&1; echo "exit $?" >> gpurun_out/b_diag.txt
done; done
echo "== tinystack bvh8 20000 rays" >> gpurun_out/b_diag.txt
VMN_LIB=$V/libvermilion_b200_tinystack.so VMN_DEBUG_NO_DEPTH_CHECK=1 timeout 60 python tools/diag_tinystack.py 8 20000 >> gpurun_out/b_diag.txt 2>&1; echo "exit $?" >> gpurun_out/b_diag.txt
echo "== tinystack bvh8 20000 rays level 5" >> gpurun_out/b_diag.txt
VMN_LIB=$V/libvermilion_b200_tinystack.so VMN_DEBUG_NO_DEPTH_CHECK=1 timeout 60 python tools/diag_tinystack.py 8 20000 5 >> gpurun_out/b_diag.txt 2>&1; echo "exit $?" >> gpurun_out/b_diag.txt
cat gpurun_out/b_diag.txt
nvidia-smi --query-gpu=name,clocks.sm --format=csv,noheader
VMN_LIB=$V/libvermilion_b200_checked.so timeout 300 python tools/sanitize_small.py > gpurun_out/b_checked.log 2>&1; echo "checked exit $?"; tail -2 gpurun_out/b_checked.log
timeout 400 python tools/ab_bench.py default default@VMN_COLLAPSE=greedy default@VMN_DP_CPRIM=0.3 default@VMN_DP_CPRIM=0.5 default@VMN_DP_CPRIM=0.7 > gpurun_out/b_ab_c2.txt 2>&1
timeout 400 python tools/ab_bench.py --c3 default default@VMN_COLLAPSE=greedy default@VMN_DP_CPRIM=0.3 default@VMN_DP_CPRIM=0.5 default@VMN_DP_CPRIM=0.7 default@VMN_REFILL=8 > gpurun_out/b_ab_c3.txt 2>&1
cat gpurun_out/b_ab_c2.txt gpurun_out/b_ab_c3.txt
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_robustness.py::test_too_deep_tree_is_refused_at_commit_and_overflow_is_reported > gpurun_out/b_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/b_gpu_tests.log
tail -3 gpurun_out/b_gpu_tests.log
timeout 300 python tools/microbench_c3.py --check 0 --reps 3 > gpurun_out/b_c3.json 2> gpurun_out/b_c3.err; echo "c3 exit $?"
python tools/c3_once.py > gpurun_out/b_c3_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_intersect -s 1 -c 1 -o gpurun_out/b_c3_intersect python tools/c3_once.py > gpurun_out/b_ncu_c3.log 2>&1; echo "ncu c3 exit $?"
python tools/c2_once.py > gpurun_out/b_c2_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade" -s 10 -c 4 -o gpurun_out/b_c2_extend_shade python tools/c2_once.py > gpurun_out/b_ncu_c2.log 2>&1; echo "ncu c2 exit $?"
ls -la gpurun_out/
