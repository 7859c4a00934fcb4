#!/bin/bash
# r02w: bench.py with its default flags (the driver's N = 1 line) and with few steps (cold e2e check), on the kernels as committed
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/w_bench_n1.json 2> gpurun_out/w_bench_n1.err; echo "bench exit $?"; cut -c1-1800 gpurun_out/w_bench_n1.json
timeout 600 python bench.py --no-cpu --no-c3 --steps 2 --warmup 1 > gpurun_out/w_bench_short.json 2> gpurun_out/w_bench_short.err; echo "short exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/w_bench_n1.json", "gpurun_out/w_bench_short.json"):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d["value"], d["e2e"], d["roofline"]["frac"], d["roofline"]["traffic"])
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/w_bench_reference.json 2> gpurun_out/w_bench_reference.err; echo "ref exit $?"; cut -c1-600 gpurun_out/w_bench_reference.json
