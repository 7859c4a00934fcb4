#!/bin/bash
# final 1-GPU call: smoke(), all GPU tests, the captures traffic.json is stamped from (gpu_call_l.sh), then bench.py with default flags
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/fin_smoke.log 2>&1; tail -2 gpurun_out/fin_smoke.log
timeout 700 python -m pytest tests -m gpu -q -x > gpurun_out/fin_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/fin_tests.log; tail -3 gpurun_out/fin_tests.log
bash tools/gpu_call_l.sh > gpurun_out/fin_l.log 2>&1; grep exit gpurun_out/fin_l.log
timeout 900 python bench.py > gpurun_out/fin_bench_n1.json 2> gpurun_out/fin_bench_n1.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference > gpurun_out/fin_bench_reference_n1.json 2> gpurun_out/fin_bench_reference_n1.err; echo "ref exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/fin_bench_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["step_ms"], d["roofline"]["frac"], d["roofline"]["traffic"], d["roofline_c3"]["mrays_bvh8"], d["roofline_c3"]["roofline"]["frac"], d["cpu_baseline"]["value"])
r = json.loads(open("gpurun_out/fin_bench_reference_n1.json").read().strip().splitlines()[-1])
print("reference", r["value"], r["ms_per_step"])
PY
