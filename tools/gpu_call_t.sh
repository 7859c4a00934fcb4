#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/t_tests.log; tail -3 gpurun_out/t_tests.log
timeout 300 python tools/ab_bench.py default default > gpurun_out/t_ab_c2.txt 2>&1; cat gpurun_out/t_ab_c2.txt
