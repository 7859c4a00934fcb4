#!/bin/bash
# after tools/gpu_call_l.sh: regenerate the r02l summaries and stamp profiles/traffic.json with the hash of the sources as they are now
set -e
cd "$(dirname "$0")/.."
python profiles/ncu_summary.py gpurun_out/l_c2_extend_shade.ncu-rep > profiles/r02l_ncu_extend_shade_c2.txt
python profiles/ncu_summary.py gpurun_out/l_c3_intersect.ncu-rep > profiles/r02l_ncu_intersect_c3.txt
cp gpurun_out/l_launches_bench_c2.csv profiles/r02l_launches_bench_c2.csv
python tools/update_traffic.py gpurun_out/l_c2_extend_shade.ncu-rep k_extend k_shade --note "C2 frame at 16 spp, 33.2 M rays in the first launch (tools/c2_once.py), three launches each (bounces 1-3)" > /dev/null
python tools/update_traffic.py gpurun_out/l_c3_intersect.ncu-rep k_intersect --note "C3, 4.19 M rays (tools/c3_once.py)" > /dev/null
python - <<'PY'
import json
p = 'profiles/traffic.json'; d = json.load(open(p))
d['k_extend']['rays_per_launch'] = 33177600; d['k_extend']['dram_bytes_per_ray'] = d['k_extend']['dram_bytes_per_launch'] / 33177600
d['k_intersect']['rays_per_launch'] = 4194304
json.dump(d, open(p, 'w'), indent=1)
print({k: (v['source_hash'], v['dram_bytes_per_launch']) for k, v in d.items()})
PY
