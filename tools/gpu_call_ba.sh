#!/bin/bash
# k_shade: block-aggregated queue append (one atomicAdd per 128 rays) against the warp-aggregated one, 64 spp in one 132.7 M-path batch
mkdir -p gpurun_out
timeout 600 python tools/ab_bench.py --spp 64 --max-paths 134217728 default blockagg default blockagg > gpurun_out/ba_ab_c2.txt 2>&1; cat gpurun_out/ba_ab_c2.txt
