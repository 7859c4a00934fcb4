#!/bin/bash
# final captures of the round: ncu --set full of k_extend / k_shade (C2) and k_intersect (C3) on the kernels as committed, launch list of bench.py
mkdir -p gpurun_out
python tools/c2_once.py > gpurun_out/l_c2_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade" -s 10 -c 6 -o gpurun_out/l_c2_extend_shade python tools/c2_once.py > gpurun_out/l_ncu_c2.log 2>&1; echo "ncu c2 exit $?"
python tools/c3_once.py > gpurun_out/l_c3_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_intersect -s 1 -c 1 -o gpurun_out/l_c3_intersect python tools/c3_once.py > gpurun_out/l_ncu_c3.log 2>&1; echo "ncu c3 exit $?"
python bench.py --no-cpu --no-c3 --steps 2 --warmup 1 > gpurun_out/l_bench_plain.json 2> gpurun_out/l_bench_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/l_launches_bench_c2.csv python bench.py --no-cpu --no-c3 --steps 2 --warmup 1 > gpurun_out/l_ncu_bench.log 2>&1; echo "launch list exit $?"
ls -la gpurun_out/l_*
