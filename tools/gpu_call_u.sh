#!/bin/bash
# r02u: last-segment attribute skip + sphere heads in the kernel parameters (A/B against the LDG variant), occupancy variants on C3 / city
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/u_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/u_tests.log; tail -3 gpurun_out/u_tests.log
timeout 400 python tools/ab_bench.py default sphldg default sphldg > gpurun_out/u_ab_c2.txt 2>&1; cat gpurun_out/u_ab_c2.txt
timeout 600 python tools/ab_bench.py --c3 default occ9 occ10 occ12 > gpurun_out/u_ab_c3.txt 2>&1; cat gpurun_out/u_ab_c3.txt
V=vermilion_b200/variants
for v in "" occ9 occ10; do
  if [ -n "$v" ]; then export VMN_LIB=$V/libvermilion_b200_$v.so; else unset VMN_LIB; fi
  echo "== ${v:-default}" >> gpurun_out/u_ab_city.txt
  timeout 300 python tools/ab_city.py 2>&1 | grep '"counters": false' >> gpurun_out/u_ab_city.txt
done
cat gpurun_out/u_ab_city.txt
