#!/bin/bash
# GPU call G (2 GPUs): the vmn_dist layer over real NCCL — tests, headless --gpus 2, bench.py --gpus 2 under torchrun (weak line, strong, C4, C5)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 600 python -m pytest tests/test_gpu_dist.py tests/test_adapter.py tests/test_headless.py tests/test_gpu_robustness.py -m gpu -q > gpurun_out/g_gpu_tests_2gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/g_gpu_tests_2gpu.log
tail -8 gpurun_out/g_gpu_tests_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/g_bench_n2.json 2> gpurun_out/g_bench_n2.err; echo "bench n2 exit $?"
tail -c 1500 gpurun_out/g_bench_n2.err
cat gpurun_out/g_bench_n2.json | cut -c1-3000
