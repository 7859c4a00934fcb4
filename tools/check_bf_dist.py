#!/usr/bin/env python
"""torchrun --nproc-per-node N tools/check_bf_dist.py — BruteForceTracer tile split over NCCL: the all-reduced frame must equal the
single-GPU frame bit for bit (every rank checks), C2 scene at 1080p; prints one JSON line on rank 0."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vermilion_b200 import distributed as vd  # noqa: E402
from vermilion_b200 import engine, scenes  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)
cam = engine.camera_from_settings(1920, 1080, (-900.0, 700.0, 1800.0), (-10.0, 25.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), 128)
with engine.Engine(device=local, stream=stream.cuda_stream) as eng:
    eng.set_scene(pos, nrm, uv)
    eng.set_camera(cam)
    eng.render(mode=engine.MODE_BRUTEFORCE, seed=1, shadow=True)
    full = eng.read_image()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(3):
        if i == 2:
            ev0.record(stream)
        img = vd.render_bruteforce_distributed(eng, torch.device("cuda", local), 1, rank, world, tile=64, shadow=True)
    ev1.record(stream)
    ev1.synchronize()
ok = bool(np.array_equal(img.view(np.uint32), full.view(np.uint32)))
t = torch.tensor([ev0.elapsed_time(ev1), float(ok)], device="cuda", dtype=torch.float64)
if world > 1:
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tmin = t.clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    ms, ok_all = float(tmax[0]), bool(tmin[1] > 0)
    dist.destroy_process_group()
else:
    ms, ok_all = float(t[0]), ok
if rank == 0:
    print(json.dumps({"check": "BruteForce tile split + NCCL all-reduce of both film planes == single-GPU frame", "n_gpus": world,
                      "bit_identical_on_every_rank": ok_all, "ms_render_reduce_readback": ms}), file=sys.stderr if False else sys.stdout)
sys.exit(0 if ok_all else 1)
