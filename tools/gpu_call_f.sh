#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/f_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/f_gpu_tests.log
tail -8 gpurun_out/f_gpu_tests.log
timeout 300 python tools/bench_bruteforce.py > gpurun_out/f_bf.json 2> gpurun_out/f_bf.err; cat gpurun_out/f_bf.json
timeout 600 python bench.py > gpurun_out/f_bench_n1.json 2> gpurun_out/f_bench_n1.err; echo "bench exit $?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; tail -2 gpurun_out/f_smoke.log
