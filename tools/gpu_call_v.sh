#!/bin/bash
# r02v: k_intersect at 9 CTAs per SM for out-of-cache scenes: event-timed C3 with and without, then the GPU tests
mkdir -p gpurun_out
VMN_OOC_OCC=0 timeout 600 python tools/microbench_c3.py --no-bvh2 --check 0 > gpurun_out/v_c3_occ8.json 2> gpurun_out/v_c3_occ8.err; cat gpurun_out/v_c3_occ8.json
timeout 600 python tools/microbench_c3.py > gpurun_out/v_c3_occ9.json 2> gpurun_out/v_c3_occ9.err; cat gpurun_out/v_c3_occ9.json
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/v_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/v_tests.log; tail -3 gpurun_out/v_tests.log
