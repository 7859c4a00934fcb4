#!/bin/bash
mkdir -p gpurun_out
timeout 400 python tools/ab_bench.py default deferonly ooconly noall default > gpurun_out/e_ab_c2.txt 2>&1
timeout 400 python tools/ab_bench.py --c3 default@VMN_OOC=0 deferonly ooconly@VMN_OOC=0 noall default@VMN_OOC=1 default@VMN_OOC=0,VMN_REFILL=8 deferonly@VMN_REFILL=2 > gpurun_out/e_ab_c3.txt 2>&1
cat gpurun_out/e_ab_c2.txt gpurun_out/e_ab_c3.txt
