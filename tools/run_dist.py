#!/usr/bin/env python
"""C4 / C5 of BASELINE.json — the multi-GPU configurations. One process per GPU (torchrun), scene replicated, only the film
crosses NVLink (vermilion_b200/distributed.py). Rank 0 prints one JSON line.

  torchrun --nproc-per-node N tools/run_dist.py --config c4 [--tess 24] [--res 3840x2160] [--spp 256]
      city (~50 M triangles), 4K, 256 spp, FIXED mode, interleaved 64x64 tiles dealt to the ranks, one all-reduce at the end.
      Full-size parity property: rank r re-renders the tiles of rank (r+1)%N and must reproduce that part of the reduced film
      bit-exactly (the render is a pure function of (pixel, sample), independent of the rank that runs it).
  torchrun --nproc-per-node N tools/run_dist.py --config c5 [--res 3840x2160] [--spp 4096] [--k 64]
      C2 scene, 4K, 4096 spp split into N sample slices, progressive: every rank renders K spp at a time into its own film and the
      films are all-reduced (into a scratch copy) at the checkpoints 64, 256, 1024, 4096 total spp. RMSE of sampled rows against
      the CPU oracle (reference code, same frame, 4096 spp, other RNG) at every checkpoint, beside the oracle-vs-oracle floor.
"""
import argparse
import json
import os

# torchrun exports OMP_NUM_THREADS=1 to every rank; rank 0's CPU-oracle leg must use all host cores (libgomp reads this when it loads).
# (The first 8-GPU run of C5 spent 11 minutes in a single-threaded oracle because of this.)
if os.environ.get("OMP_NUM_THREADS") == "1" and os.environ.get("RANK", "0") == "0":
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["c4", "c5"])
    ap.add_argument("--res", default="3840x2160")
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--tess", type=int, default=24)
    ap.add_argument("--k", type=int, default=64)
    ap.add_argument("--tile", type=int, default=64)
    ap.add_argument("--cpu-rows", type=int, default=8)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--max-paths", type=int, default=16 << 20)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from vermilion_b200 import distributed as vd
    from vermilion_b200 import engine, scenes

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    width, height = (int(x) for x in args.res.lower().split("x"))
    npix = width * height

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def allsum(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t[0])

    t0 = time.time()
    if args.config == "c4":
        spp = args.spp or 256
        pos, nrm, uv = scenes.city(64, 64, args.tess, 2024)
        cam_pos, cam_rot = (0.0, 900.0, 1950.0), (25.0, 0.0, 0.0)
        workload = f"C4: city 64x64 blocks, {len(pos)} triangles, {width}x{height}, {spp} spp, FIXED, interleaved {args.tile}x{args.tile} tiles over {world} GPU(s)"
    else:
        spp = args.spp or 4096
        pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)
        cam_pos, cam_rot = (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0)
        workload = f"C5: C2 scene ({len(pos)} triangles), {width}x{height}, {spp} spp progressive, sample slices over {world} GPU(s), all-reduce at checkpoints"
    gen_s = time.time() - t0
    eng = engine.Engine(device=local, max_paths=args.max_paths, stream=stream.cuda_stream)
    eng.set_scene(pos, nrm, uv)
    build_ms = eng.stats()["build_ms"]
    cam = engine.camera_from_settings(width, height, cam_pos, cam_rot, 6.0, (3.6, 3.6 * height / width), spp)
    eng.set_camera(cam)
    film = vd.device_film_tensor(eng, dev)
    out = {"config": args.config, "workload": workload, "n_gpus": world, "scene_gen_s": gen_s, "bvh_build_ms": build_ms}

    if args.config == "c4":
        share = vd.render_params("tile", rank, world, spp, tile=args.tile)
        # warm-up: a 4-spp frame of the same share
        eng.set_camera(engine.camera_from_settings(width, height, cam_pos, cam_rot, 6.0, (3.6, 3.6 * height / width), 4))
        eng.render(mode=engine.MODE_FIXED, seed=1, **share)
        eng.set_camera(cam)
        sync()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        eng.render(mode=engine.MODE_FIXED, seed=1, **share)
        st = eng.stats()
        vd.reduce_film(film)
        b.record(stream)
        b.synchronize()
        sync()
        ms = allmax(a.elapsed_time(b))
        rays = allsum(st["rays_traced"])
        reduced = film.clone()
        out.update(ms=ms, mrays_per_s=rays / ms / 1e3, rays=rays, spp_per_s=spp / (ms / 1e3), render_ms_rank0=st["total_ms"], extend_ms_rank0=st["extend_ms"],
                   shade_ms_rank0=st["shade_ms"], spp_per_s_1080p_equiv=spp * npix / (1920 * 1080) / (ms / 1e3))
        # parity property at full size: re-render the next rank's tiles here, compare with the reduced film bit for bit
        other = (rank + 1) % world
        eng.render(mode=engine.MODE_FIXED, seed=1, **vd.render_params("tile", other, world, spp, tile=args.tile))
        torch.cuda.synchronize()
        mine = film.view(-1, 4)
        own = torch.from_numpy(vd.tile_owner_map(width, height, args.tile, world).reshape(-1) == other).to(dev)
        same = bool(torch.equal(mine[own], reduced.view(-1, 4)[own]))
        covered = bool((reduced.view(-1, 4)[:, 3] == spp).all())
        out["tiles_rerendered_by_other_rank_bit_exact"] = bool(allsum(0.0 if same else 1.0) == 0.0)
        out["every_pixel_has_spp_samples"] = covered
        img = eng.resolve_device(reduced.data_ptr())
        out["image_mean_rgb"] = [float(x) for x in img[:, :3].mean(0)]
    else:
        k = args.k
        b0, e0 = vd.sample_slice(rank, world, spp)
        checkpoints = [c for c in (64, 256, 1024, 4096, spp) if c <= spp]
        checkpoints = sorted(set(checkpoints))
        scratch = torch.empty_like(film)
        rows = [int((i + 0.5) * height / args.cpu_rows) for i in range(args.cpu_rows)]
        images = {}
        eng.render(mode=engine.MODE_FIXED, seed=1, sample_begin=b0, sample_end=min(e0, b0 + 1))  # warm-up (film cleared again below)
        sync()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        done_local, rays, first, reduce_ms = 0, 0, True, 0.0
        for cp in checkpoints:
            want = (e0 - b0) * cp // spp  # local samples done when the job as a whole has cp
            while done_local < want:
                n = min(k, want - done_local)
                eng.render(mode=engine.MODE_FIXED, seed=1, sample_begin=b0 + done_local, sample_end=b0 + done_local + n, accumulate=not first)
                first = False
                rays += eng.stats()["rays_traced"]
                done_local += n
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            r0.record(stream)
            scratch.copy_(film)
            vd.reduce_film(scratch)
            r1.record(stream)
            r1.synchronize()
            reduce_ms += r0.elapsed_time(r1)
            if rank == 0:
                s = scratch.view(height, width, 4)[rows].reshape(-1, 4).cpu().numpy()
                images[cp] = vd.resolve(s)
        b.record(stream)
        b.synchronize()
        sync()
        ms = allmax(a.elapsed_time(b))
        rays = allsum(rays)
        out.update(ms=ms, mrays_per_s=rays / ms / 1e3, rays=rays, spp_per_s=spp / (ms / 1e3), spp_per_s_1080p_equiv=spp * npix / (1920 * 1080) / (ms / 1e3),
                   allreduce_ms_total=reduce_ms, allreduce_count=len(checkpoints), film_bytes=int(film.numel() * 4), k=k, checkpoints=checkpoints)
        total_n = float(scratch.view(-1, 4)[:, 3].min()), float(scratch.view(-1, 4)[:, 3].max())
        out["samples_per_pixel_min_max"] = total_n
        if rank == 0 and not args.no_cpu:
            import checker
            cs = checker.camera_settings(width, height, cam_pos, cam_rot, 6.0, (3.6, 3.6 * height / width), spp)
            ocam = checker.camera_fields_py(cs)
            use_ref = checker.Ref.available("ieee")
            t0 = time.time()
            cpu = []
            for seed in (11, 12):
                parts = []
                if use_ref:
                    ref = checker.Ref(pos, nrm, uv)
                    for r in rows:
                        parts.append(ref.render_seeded(cs, mode=checker.MODE_FIXED, seed=seed, p0=r * width, p1=(r + 1) * width)[r * width:(r + 1) * width])
                else:
                    orc = checker.Oracle(pos, nrm, uv)
                    for r in rows:
                        parts.append(orc.render(ocam, checker.MODE_FIXED, checker.RNG_MT64, seed, r * width, (r + 1) * width)[0][r * width:(r + 1) * width])
                cpu.append(np.concatenate(parts))
            cpu_s = time.time() - t0
            rm = lambda x, y: float(np.sqrt(np.mean((x[:, :3] - y[:, :3]) ** 2)))
            out["rmse"] = {"rows": rows, "cpu_kind": "reference (oracle/_ref, -O2 -ffp-contract=off)" if use_ref else "oracle port", "cpu_spp": spp, "cpu_seconds": cpu_s,
                           "cpu_vs_cpu_floor": rm(cpu[0], cpu[1]), "gpu_vs_cpu_by_checkpoint": {str(cp): rm(images[cp], cpu[0]) for cp in checkpoints},
                           "bar": "gpu_vs_cpu at the final checkpoint <= 2 x cpu_vs_cpu_floor (SURVEY 8c)"}
            out["rmse"]["pass"] = out["rmse"]["gpu_vs_cpu_by_checkpoint"][str(checkpoints[-1])] <= 2 * out["rmse"]["cpu_vs_cpu_floor"]
    if rank == 0:
        print(json.dumps(out))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
