#!/bin/bash
# GPU call D: bisect the k_extend regression of call C (deferred store / abandon-on-overflow / OOC code)
mkdir -p gpurun_out
timeout 400 python tools/ab_bench.py default nodefer noabandon noooc noall > gpurun_out/d_ab_c2.txt 2>&1
timeout 400 python tools/ab_bench.py --c3 default@VMN_OOC=0 nodefer@VMN_OOC=0 noabandon@VMN_OOC=0 noooc noall default@VMN_OOC=1 nodefer@VMN_OOC=1 > gpurun_out/d_ab_c3.txt 2>&1
cat gpurun_out/d_ab_c2.txt gpurun_out/d_ab_c3.txt
