#!/bin/bash
# last 1-GPU call of round 2 (re-created container, libraries rebuilt from the committed sources): smoke(), all GPU tests, bench.py
mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/fin2_smoke.log 2>&1; tail -2 gpurun_out/fin2_smoke.log
timeout 300 python -m pytest tests -m gpu -q -x > gpurun_out/fin2_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/fin2_tests.log; tail -3 gpurun_out/fin2_tests.log
timeout 240 python bench.py > gpurun_out/fin2_bench_n1.json 2> gpurun_out/fin2_bench_n1.err; echo "bench exit $?"
tail -c 600 gpurun_out/fin2_bench_n1.json
