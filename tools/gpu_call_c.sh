#!/bin/bash
# GPU call C of round 2: overflow diagnostics (abandon-on-overflow), full GPU test suite, OOC prefetch / deferred-store A/B, coherence bound,
# one-launch frames (C1, BruteForce), bench line
mkdir -p gpurun_out
V=vermilion_b200/variants
: > gpurun_out/c_diag.txt
for spec in "tinystack 8" "tinystack 2" "sh2 8"; do set -- $spec
  echo "== $1 bvh$2 20000 rays" >> gpurun_out/c_diag.txt
  VMN_LIB=$V/libvermilion_b200_$1.so VMN_DEBUG_NO_DEPTH_CHECK=1 timeout 40 python tools/diag_tinystack.py $2 20000 >> gpurun_out/c_diag.txt 2>&1; echo "exit $?" >> gpurun_out/c_diag.txt
done
cat gpurun_out/c_diag.txt
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_robustness.py::test_too_deep_tree_is_refused_at_commit_and_overflow_is_reported > gpurun_out/c_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/c_gpu_tests.log
tail -15 gpurun_out/c_gpu_tests.log
timeout 500 python tools/ab_bench.py --c3 default default@VMN_OOC=0 default@VMN_OOC=1 default@VMN_OOC=2 default@VMN_L2_PERSIST=2 default@VMN_OOC=0,VMN_L2_PERSIST=2 default@VMN_REFILL=8 default@VMN_REFILL=2 > gpurun_out/c_ab_c3.txt 2>&1
timeout 300 python tools/ab_bench.py default default@VMN_OOC=3 default@VMN_REFILL=4 > gpurun_out/c_ab_c2.txt 2>&1
cat gpurun_out/c_ab_c3.txt gpurun_out/c_ab_c2.txt
timeout 300 python tools/ab_coherence.py --scene c2 > gpurun_out/c_coherence_c2.json 2> gpurun_out/c_coherence_c2.err; echo "coh c2 $?"
timeout 400 python tools/ab_coherence.py --scene city:32 > gpurun_out/c_coherence_city32.json 2> gpurun_out/c_coherence_city.err; echo "coh city $?"
cat gpurun_out/c_coherence_c2.json gpurun_out/c_coherence_city32.json
timeout 300 python tools/bench_c1.py > gpurun_out/c_c1.json 2> gpurun_out/c_c1.err; echo "c1 $?"; cat gpurun_out/c_c1.json
VMN_MEGA=0 timeout 300 python tools/bench_c1.py > gpurun_out/c_c1_wavefront.json 2>> gpurun_out/c_c1.err; cat gpurun_out/c_c1_wavefront.json
timeout 300 python tools/bench_bruteforce.py > gpurun_out/c_bf.json 2> gpurun_out/c_bf.err; echo "bf $?"; cat gpurun_out/c_bf.json
VMN_MEGA=0 timeout 300 python tools/bench_bruteforce.py > gpurun_out/c_bf_wavefront.json 2>> gpurun_out/c_bf.err; cat gpurun_out/c_bf_wavefront.json
timeout 600 python bench.py > gpurun_out/c_bench_n1.json 2> gpurun_out/c_bench_n1.err; echo "bench exit $?"
