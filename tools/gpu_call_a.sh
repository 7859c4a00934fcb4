#!/bin/bash
# GPU call A of round 2: full GPU test suite, bench line with roofline_c3, A/B of the cache-policy / deferral knobs, sanitizer passes.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/a_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/a_gpu_tests.log
tail -3 gpurun_out/a_gpu_tests.log
timeout 600 python bench.py > gpurun_out/a_bench_n1.json 2> gpurun_out/a_bench_n1.err; echo "bench exit $?"
timeout 400 python tools/ab_bench.py default nohints default@VMN_L2_PERSIST=0 nohints@VMN_L2_PERSIST=0 defer24 defer32 deferinf default@VMN_REFILL=4 default@VMN_REFILL=12 deferinf@VMN_REFILL=4 deferinf@VMN_REFILL=12 defer32@VMN_REFILL=12 > gpurun_out/a_ab_c2.txt 2>&1
timeout 400 python tools/ab_bench.py --c3 default nohints defer32 deferinf default@VMN_REFILL=4 default@VMN_REFILL=16 > gpurun_out/a_ab_c3.txt 2>&1
cat gpurun_out/a_ab_c2.txt gpurun_out/a_ab_c3.txt
timeout 420 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_small.py > gpurun_out/a_sanitizer_memcheck.log 2>&1; echo "memcheck exit $?"
timeout 420 compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitize_small.py > gpurun_out/a_sanitizer_racecheck.log 2>&1; echo "racecheck exit $?"
tail -4 gpurun_out/a_sanitizer_memcheck.log gpurun_out/a_sanitizer_racecheck.log
