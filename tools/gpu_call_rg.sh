#!/bin/bash
# raygen: 32-bit slot divide + direct (atomic-free) slots for contiguous FIXED rounds; A/B against the warp-aggregated append
mkdir -p gpurun_out
timeout 400 python tools/ab_bench.py default rgatomic default rgatomic > gpurun_out/rg_ab_c2.txt 2>&1; cat gpurun_out/rg_ab_c2.txt
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/rg_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/rg_tests.log; tail -3 gpurun_out/rg_tests.log
