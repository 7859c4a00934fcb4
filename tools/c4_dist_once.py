#!/usr/bin/env python
"""C4 through a rank group (torchrun, N >= 2): scene load (upload on rank 0 + broadcast + build) and one tiled 4K frame, once per value of
VMN_LOAD_OVERLAP given on the command line — the A/B of the pipelined scene load at 52 M triangles.
  torchrun --nproc-per-node 2 tools/c4_dist_once.py [1 0 1] [--blocks 64] [--spp 64]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from vermilion_b200 import engine, scenes

ap = argparse.ArgumentParser()
ap.add_argument("values", nargs="*", default=["1", "0", "1"])
ap.add_argument("--blocks", type=int, default=64)
ap.add_argument("--spp", type=int, default=64)
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
uid = torch.zeros(engine.DIST_ID_BYTES, dtype=torch.uint8, device="cuda")
if rank == 0:
    uid.copy_(torch.tensor(list(engine.DistGroup.unique_id()), dtype=torch.uint8))
dist.broadcast(uid, 0)
grp = engine.DistGroup.rank(bytes(uid.cpu().tolist()), rank, world, device=local, max_paths=128 << 20, stream=stream.cuda_stream)
pos = nrm = uv = None
if rank == 0:
    pos, nrm, uv = scenes.city(args.blocks, args.blocks, 24, 2024)
w, h = 3840, 2160
mk = lambda spp: engine.camera_from_settings(w, h, (0.0, 900.0, 1950.0), (25.0, 0.0, 0.0), 6.0, (3.6, 3.6 * h / w), spp)
for v in args.values:
    os.environ["VMN_LOAD_OVERLAP"] = v
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    grp.set_scene(pos, nrm, uv, broadcast=True)
    torch.cuda.synchronize()
    load_s = time.perf_counter() - t0
    grp.set_camera(mk(4))
    grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=engine.DIST_SPLIT_TILES)
    grp.set_camera(mk(args.spp))
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record(stream)
    grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=engine.DIST_SPLIT_TILES)
    b.record(stream); b.synchronize()
    host_s = time.perf_counter() - t0
    ms = a.elapsed_time(b)
    st = grp.member_stats(0)
    out = {"rank": rank, "overlap": v, "triangles": st["n_triangles"], "load_s": round(load_s, 3), "scene_ms": round(grp.stats()["scene_ms"], 1),
           "build_ms": round(st["build_ms"], 1), "frame_event_ms": round(ms, 1), "frame_host_ms": round(host_s * 1e3, 1), "render_total_ms": round(st["total_ms"], 1),
           "extend_ms": round(st["extend_ms"], 1), "shade_ms": round(st["shade_ms"], 1), "collective_ms": round(grp.stats()["collective_ms"], 2),
           "mrays_rank": round(st["rays_traced"] / ms / 1e3, 1)}
    for r in range(world):
        if r == rank:
            print(json.dumps(out), flush=True)
        dist.barrier()
grp.close()
dist.destroy_process_group()
