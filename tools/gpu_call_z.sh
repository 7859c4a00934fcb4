#!/bin/bash
# r02z (2 GPUs): does the C4 wall-over-slowest-rank gap of bench.py reproduce, and does it follow VMN_LOAD_OVERLAP?
mkdir -p gpurun_out
for v in 1 0; do
VMN_LOAD_OVERLAP=$v timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$v bench.py --gpus 2 --steps 2 --warmup 3 --no-cpu --no-c3 > gpurun_out/z_bench_n2_ov$v.json 2> gpurun_out/z_bench_n2_ov$v.err; echo "bench ov=$v exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/z_bench_n2_ov$v.json").read().strip().splitlines()[-1])
print($v, d["value"], d["e2e"]["value"], d["e2e"]["step_ms"])
print(json.dumps(d["c4"])[:700])
PY
done
