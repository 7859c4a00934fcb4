#!/bin/bash
# tile-mode: power-of-two tile shifts in map_pixel (k_shade, k_raygen) on top of the direct raygen slots; then the GPU tests
mkdir -p gpurun_out
timeout 400 python tools/ab_tiles.py --spp 64 default default > gpurun_out/tl2_ab_tiles.txt 2>&1
timeout 300 python tools/ab_tiles.py --spp 64 --world 8 default default >> gpurun_out/tl2_ab_tiles.txt 2>&1
timeout 400 python tools/ab_bench.py --spp 64 --max-paths 134217728 default default >> gpurun_out/tl2_ab_tiles.txt 2>&1; cat gpurun_out/tl2_ab_tiles.txt
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/tl2_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/tl2_tests.log; tail -3 gpurun_out/tl2_tests.log
