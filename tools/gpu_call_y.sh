#!/bin/bash
# r02y (2 GPUs): A/B of the pipelined scene load on the 52 M-triangle city through a rank group
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/c4_dist_once.py 1 1 0 --spp 256 > gpurun_out/y_c4_load_ab.txt 2> gpurun_out/y_c4_load_ab.err; echo "exit $?"
cat gpurun_out/y_c4_load_ab.txt; tail -5 gpurun_out/y_c4_load_ab.err
