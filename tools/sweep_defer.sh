#!/bin/bash
# Next-round measurement: builds VMN_TRI_DEFER variants here (CPU, nvcc cross-compiles) and prints the gpurun
# command that A/Bs them, together with the refill threshold, on the 16-spp C2 frame and on C3.
#   bash tools/sweep_defer.sh            # build
#   gpurun --timeout 300 -- '<printed command>'
set -e
cd "$(dirname "$0")/../vermilion_b200/csrc"
for k in 0 8 24 32 64; do
  make VARIANT=defer$k EXTRA="-DVMN_TRI_DEFER=$k" -j4 > /dev/null
  grep -A3 "k_extendILb0" ../variants/obj_defer$k/wavefront.o.ptxas.log | grep -i "registers" | sed "s/^/defer$k: /"
done
echo "python tools/ab_bench.py default defer0 defer8 defer24 defer32 defer64 default@VMN_REFILL=4 default@VMN_REFILL=12 defer32@VMN_REFILL=4 defer32@VMN_REFILL=12 | tee gpurun_out/ab_defer_sweep.txt; python tools/ab_bench.py --c3 default defer0 defer32 defer64 | tee gpurun_out/ab_defer_sweep_c3.txt"
