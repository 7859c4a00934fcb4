#!/bin/bash
# ncu --set full of k_extend on the out-of-cache city (12 M triangles): what binds the C4 regime
mkdir -p gpurun_out
python tools/city_once.py > gpurun_out/city_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_extend" -s 5 -c 4 -o gpurun_out/city_extend python tools/city_once.py > gpurun_out/city_ncu.log 2>&1; echo "ncu city exit $?"
cat gpurun_out/city_plain.log
