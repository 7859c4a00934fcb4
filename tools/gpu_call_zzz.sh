#!/bin/bash
# r02zzz: vmn_intersect_static (static-scratch entry point). The new test alone in a fresh process, then the whole GPU suite (the new file
# is collected last), then the default bench on the rebuilt library.
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/zzz_no_alloc_alone.log 2>&1; echo "alone exit $?"; tail -3 gpurun_out/zzz_no_alloc_alone.log
timeout 200 python -m pytest tests -x -q -m gpu > gpurun_out/zzz_gpu_tests.log 2>&1; echo "suite exit $?"; tail -3 gpurun_out/zzz_gpu_tests.log
timeout 170 python bench.py > gpurun_out/zzz_bench_n1.json 2> gpurun_out/zzz_bench_n1.err; echo "bench exit $?"; tail -c 600 gpurun_out/zzz_bench_n1.json
