#!/bin/bash
# r02x: geometry build inside vmn_scene_set_triangles, beside the normals / uvs upload: GPU tests, then the e2e steps of bench.py
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/x_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/x_tests.log; tail -3 gpurun_out/x_tests.log
timeout 600 python bench.py --no-cpu --no-c3 --steps 3 --warmup 3 > gpurun_out/x_bench_short.json 2> gpurun_out/x_bench_short.err; echo "short exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/x_bench_short.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"], d["bvh_build_ms"])
PY
