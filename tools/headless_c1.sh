python - <<'PY'
import sys; sys.path.insert(0,'.')
from vermilion_b200 import scenes
pos,nrm,uv=scenes.cornell()
mats=[("white",(.73,.73,.73),0,6),("red",(.65,.05,.05),6,8),("green",(.12,.45,.15),8,10),("blocks",(.73,.73,.73),10,30)]
scenes.write_obj('/tmp/cornell.obj',pos,nrm,uv,mats)
PY
H=oracle/_ref/vermilion_headless
A="--res 512x512 --spp 16 --camera 0,275,1100,0,0,0 --film 6,3.6,3.6 --quiet"
for e in cpu b200 b200 cpu-bf b200-bf b200-bf; do $H /tmp/cornell.obj $A --engine $e --out gpurun_out/r01ah_c1_$e | grep '^{' ; done > gpurun_out/r01ah_headless_c1.jsonl 2>&1
cat gpurun_out/r01ah_headless_c1.jsonl
timeout 330 python -m pytest tests -m gpu -q > gpurun_out/r01ah_gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r01ah_gpu_tests.log; tail -4 gpurun_out/r01ah_gpu_tests.log
