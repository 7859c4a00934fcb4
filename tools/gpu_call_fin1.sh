#!/bin/bash
# final 1-GPU call: all GPU tests, then the captures traffic.json is stamped from (gpu_call_l.sh), then bench.py with default flags
mkdir -p gpurun_out
timeout 700 python -m pytest tests -m gpu -q -x > gpurun_out/fin_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/fin_tests.log; tail -3 gpurun_out/fin_tests.log
bash tools/gpu_call_l.sh
