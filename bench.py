#!/usr/bin/env python
"""bench.py — Mrays/s of the B200 path-tracing hot path on BASELINE.json's configs[1] (C2):
procedural tessellated-sphere scene, 983 040 triangles, 1920x1080, 64 spp, primary + 4 bounces (FIXED mode).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--spp S] [--res WxH]

A step = one full frame through the wavefront (raygen -> 5 x (extend, shade) -> accumulate per sample batch).
N > 1 (torchrun, one rank per GPU): sample-slice split — every rank renders the full frame for its own 64-sample
slice of an N*64-spp frame (weak scaling); the film sums are reduced to rank 0 over NCCL inside the timed region by the
library's own multi-GPU layer (vmn_dist_*, csrc/dist.cu: the same C++ path the Integrator adapter uses). The N > 1 line also
carries `strong`: the SAME 64-spp C2 frame split over the N GPUs (interleaved tiles + gather, and sample slices + reduce).
`value`   device-timed, scene resident in HBM.            `e2e`  through the C ABI with host buffers: scene upload
+ GPU BVH build + render + image read-back every step.    `roofline`  for the dominant kernel (k_extend).
`cpu_baseline` / --impl reference: the reference's own CPU code (oracle/_ref, -Ofast as shipped) on the host cores.
"""
import argparse
import ctypes
import json
import os

# torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU reference arm must use all host cores (libgomp reads this when it loads)
if os.environ.get("OMP_NUM_THREADS") == "1" and "--impl" in " ".join(os.sys.argv):
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "C2: 12 icospheres x 81920 = 983040 triangles in the reference room, 1920x1080, 64 spp, FIXED mode, <=5 segments/path"
CAM_POS, CAM_ROT = (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def make_scene():
    from vermilion_b200 import scenes
    return scenes.spheres(4, 3, 6, 150.0)


def make_camera(width, height, spp):
    from vermilion_b200 import engine
    return engine.camera_from_settings(width, height, CAM_POS, CAM_ROT, 6.0, (3.6, 3.6 * height / width), spp)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- CPU reference arm ----------------------------------------------------------------------------------------
def _bit_reversed(n):
    """0..n-1 in bit-reversed order: any prefix of it is spread evenly over the range."""
    bits = max(1, (n - 1).bit_length())
    order = [int(format(i, f"0{bits}b")[::-1], 2) for i in range(1 << bits)]
    return [i for i in order if i < n]


def cpu_reference_sample(pos, nrm, uv, width, height, spp, budget_s=12.0, threads=0, rows_per_band=8):
    """Times the reference's own CPU code (oracle/_ref/libvermref_ofast.so: unmodified Triangle/BVH/MeshEngine::RayCast, -Ofast as
    shipped) on a bounded sample of the workload: 8-row bands taken in bit-reversed order (so any prefix is spread over the frame),
    all `spp` samples, FIXED mode, until `budget_s` seconds of CPU work are spent. Rays are counted by the bit-identical oracle port on
    the same pixels (untimed; the reference has no ray counter). Falls back to timing the port if oracle/_ref is absent."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import checker
    cs = checker.camera_settings(width, height, CAM_POS, CAM_ROT, 6.0, (3.6, 3.6 * height / width), spp)
    kind = "reference" if checker.Ref.available("ofast") else "port"
    orc = checker.Oracle(pos, nrm, uv)
    ocam = checker.camera_fields_py(cs)
    t_build0 = time.time()
    ref = checker.Ref(pos, nrm, uv, variant="ofast") if kind == "reference" else None
    build_s = time.time() - t_build0
    cores = checker.Oracle.max_threads() if threads <= 0 else threads
    nbands = -(-height // rows_per_band)
    secs, rays, paths, used = 0.0, 0, 0, 0
    for band in _bit_reversed(nbands):
        r0, r1 = band * rows_per_band, min(height, (band + 1) * rows_per_band)
        p0, p1 = r0 * width, r1 * width
        t0 = time.time()
        if ref is not None:
            ref.render_seeded(cs, mode=checker.MODE_FIXED, seed=1, p0=p0, p1=p1, s0=0, s1=spp, max_segments=5, threads=threads)
        else:
            orc.render(ocam, checker.MODE_FIXED, checker.RNG_MT64, 1, p0, p1, 0, spp, 5, threads)
        secs += time.time() - t0
        _, n = orc.render(ocam, checker.MODE_FIXED, checker.RNG_MT64, 1, p0, p1, 0, spp, 5, 0)
        rays += n
        paths += (p1 - p0) * spp
        used += 1
        if secs > budget_s:
            break
    return {"value": rays / secs / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind,
            "sample": f"{used} of {nbands} {rows_per_band}-row bands (bit-reversed order, spread over the {width}x{height} frame), all {spp} spp, FIXED mode: "
                      f"{paths} paths, {rays} rays in {secs:.1f} s of CPU time (reference BVH build {build_s:.1f} s not included)",
            "spp_per_s": paths / secs / (width * height)}


def run_reference(args, width, height):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pos, nrm, uv = make_scene()
    vals = []
    t_all = time.time()
    for i in range(args.warmup + args.steps):
        r = cpu_reference_sample(pos, nrm, uv, width, height, args.spp, budget_s=args.cpu_budget / max(1, args.warmup + args.steps))
        if i >= args.warmup:
            vals.append(r)
    v = float(np.mean([x["value"] for x in vals]))
    last = vals[-1]
    last["value"] = v
    print(json.dumps({"impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": (time.time() - t_all) * 1000 / (args.warmup + args.steps), "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": WORKLOAD, "resolution": f"{width}x{height}", "spp": args.spp}, "cpu_baseline": last,
                      "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ---- GPU arm --------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--res", default="1920x1080")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-c4c5", action="store_true", help="skip the C4 / C5 records of an N > 1 run (52 M-triangle city at 4K; progressive 4K render)")
    ap.add_argument("--no-c3", action="store_true", help="skip the roofline_c3 record (10 M-triangle soup, N = 1 only)")
    ap.add_argument("--max-paths", type=int, default=128 << 20, help="paths in flight (112 B each = 15 GB at the default: the whole 64-spp 1080p frame is ONE batch); larger batches amortise the tail of every persistent launch: 16 Mi 6826, 32 Mi 7063, 64 Mi 7172, 128 Mi 7221 Mrays/s (profiles/r02j_maxpaths.txt)")
    args = ap.parse_args()
    width, height = (int(x) for x in args.res.lower().split("x"))
    if args.impl == "reference":
        return run_reference(args, width, height)

    import torch
    import torch.distributed as dist
    from vermilion_b200 import engine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: libraries that write there (NCCL prints its version line to stdout when NCCL_DEBUG is set)
    # are sent to stderr for the duration of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)

    pos, nrm, uv = make_scene()
    spp_total = args.spp * world
    cam = make_camera(width, height, spp_total)
    npix = width * height
    dev = torch.device("cuda", local)

    def new_engine(counters=False):
        e = engine.Engine(device=local, counters=counters, max_paths=args.max_paths, stream=stream.cuda_stream)
        return e

    # N = 1: one context. N > 1: a rank group of the library's multi-GPU layer (one member per process; rank 0's NCCL id is handed to
    # the others over torch.distributed, which is used for nothing else on the data path)
    grp = None
    if world > 1:
        uid = torch.zeros(engine.DIST_ID_BYTES, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid.copy_(torch.tensor(list(engine.DistGroup.unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        grp = engine.DistGroup.rank(bytes(uid.cpu().tolist()), rank, world, device=local, max_paths=args.max_paths, stream=stream.cuda_stream)
        grp.set_scene(pos, nrm, uv)
        grp.set_camera(cam)
        eng = None
    else:
        eng = new_engine()
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(cam)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def step():
        if grp is not None:
            grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=engine.DIST_SPLIT_SAMPLES)  # 64 spp per rank + ncclReduce to rank 0
        else:
            eng.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)

    def step_stats():
        return grp.member_stats(0) if grp is not None else eng.stats()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        """`steps` calls of fn, each bracketed by CUDA events on the launching stream, L2 flushed before each; returns (sum of ms, MAX
        over ranks) and the per-step stats"""
        for _ in range(warmup):
            fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        stats = []
        for a, b in ev:
            flush.zero_()
            if world > 1:
                dist.barrier()  # every rank starts the step together: the collective inside must not absorb a late starter's skew
            a.record(stream)
            fn()
            b.record(stream)
            b.synchronize()
            stats.append(step_stats())
        barrier()
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), stats

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    ms, sts = timed(step, args.steps, 0)
    clocks = sampler.stop() if sampler else None
    rays = sum(st["rays_traced"] for st in sts)
    extend_ms = sum(st["extend_ms"] for st in sts)
    launches = sum(st["kernel_launches"] for st in sts)
    t = torch.tensor([float(rays)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    rays_all = float(t[0])
    value = rays_all / (ms / 1e3) / 1e6
    paths_all = float(npix) * args.spp * world * args.steps

    # ---- strong scaling (N > 1): the SAME 64-spp frame over N GPUs, through the same group ----
    strong = None
    if grp is not None:
        grp.set_camera(make_camera(width, height, args.spp))
        strong = {"workload": f"the {args.spp}-spp C2 frame split over {world} GPUs (total work fixed)", "single_gpu_ms_per_frame": ms / args.steps,
                  "single_gpu_note": "= ms_per_step of the weak line: there every rank renders this whole frame"}
        for name, split in (("tiles", engine.DIST_SPLIT_TILES), ("samples", engine.DIST_SPLIT_SAMPLES)):
            fn = lambda: grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=split)
            sms, ssts = timed(fn, args.steps, args.warmup)
            mine = torch.tensor([sum(st["total_ms"] for st in ssts) / args.steps], dtype=torch.float64, device=dev)
            allr = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allr, mine)
            per_rank = [float(x[0]) for x in allr]
            coll = grp.stats()["collective_ms"]
            strong[name] = {"ms_per_frame": sms / args.steps, "efficiency": (ms / args.steps) / (world * (sms / args.steps)),
                            "render_ms_per_rank": per_rank, "slowest_rank_ms": max(per_rank), "imbalance": max(per_rank) / (sum(per_rank) / world),
                            "assembly_ms_root_last_frame": coll,
                            "exchange": "ncclSend/ncclRecv of each rank's own tiles (1/N of the film) to rank 0" if name == "tiles" else "ncclReduce(sum) of the film to rank 0"}
        grp.set_camera(cam)

    # ---- C4 / C5 of BASELINE.json (N > 1 only): the 52 M-triangle city at 4K by tiles, and a progressive 4K render with the film
    # assembled at checkpoints — through the same group, so that the driver's scaling run exercises them
    c4 = c5 = None
    if grp is not None and not args.no_c4c5:
        from vermilion_b200 import scenes as _scenes
        try:
            t0 = time.time()
            root = rank == 0
            cpos = cnrm = cuv = None
            if root:
                cpos, cnrm, cuv = _scenes.city(64, 64, 24, 2024)
            gen_s = time.time() - t0
            grp.set_scene(cpos, cnrm, cuv, broadcast=True)  # 5 GB up on rank 0's link, then NVLink
            ntri = grp.member_stats(0)["n_triangles"]
            del cpos, cnrm, cuv
            c4w, c4h, c4spp = 3840, 2160, 256
            mk = lambda spp: engine.camera_from_settings(c4w, c4h, (0.0, 900.0, 1950.0), (25.0, 0.0, 0.0), 6.0, (3.6, 3.6 * c4h / c4w), spp)
            grp.set_camera(mk(4))
            grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=engine.DIST_SPLIT_TILES)  # warm-up
            grp.set_camera(mk(c4spp))
            fn = lambda: grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=engine.DIST_SPLIT_TILES)
            cms, csts = timed(fn, 1, 0)
            crays = torch.tensor([float(csts[0]["rays_traced"]), csts[0]["total_ms"]], dtype=torch.float64, device=dev)
            allr = [torch.zeros_like(crays) for _ in range(world)]
            dist.all_gather(allr, crays)
            tot_rays = sum(float(x[0]) for x in allr)
            per_rank = [float(x[1]) for x in allr]
            full = None
            if root:
                film = grp.read_image(engine.FILM_SUM)
                full = bool((film[:, 3] == c4spp).all())
                del film
            c4 = {"workload": f"C4: city 64x64 blocks, {ntri} triangles, {c4w}x{c4h}, {c4spp} spp, FIXED, interleaved 64x64 tiles over {world} GPUs, tile gather to rank 0",
                  "ms_per_frame": cms, "mrays_per_s": tot_rays / cms / 1e3, "spp_per_s": c4spp / (cms / 1e3), "render_ms_per_rank": per_rank,
                  "slowest_rank_ms": max(per_rank), "wall_over_slowest_rank": cms / max(per_rank), "assembly_ms_root": grp.stats()["collective_ms"],
                  "every_pixel_has_spp_samples": full, "scene_gen_s_rank0": gen_s, "bvh_build_ms": grp.member_stats(0)["build_ms"]}
        except Exception as e:
            c4 = {"error": repr(e)}
        try:
            grp.set_scene(pos, nrm, uv)  # back to C2
            c5w, c5h, per_gpu, k = 3840, 2160, 512, 64
            c5spp = per_gpu * world
            grp.set_camera(engine.camera_from_settings(c5w, c5h, CAM_POS, CAM_ROT, 6.0, (3.6, 3.6 * c5h / c5w), c5spp))
            checkpoints = [c for c in (64 * world, 256 * world, c5spp) if c <= c5spp]
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if world > 1:
                dist.barrier()
            a.record(stream)
            done, rays5, asm_ms, first = 0, 0, 0.0, True
            for cp in checkpoints:
                while done < cp:
                    step_n = min(k * world, cp - done)
                    last = done + step_n == cp
                    grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=engine.DIST_SPLIT_SAMPLES, sample_begin=done,
                               sample_end=done + step_n, accumulate=not first, assemble=last)
                    first = False
                    done += step_n
                    rays5 += grp.member_stats(0)["rays_traced"]
                    if last:
                        asm_ms += grp.stats()["collective_ms"]
            b.record(stream); b.synchronize()
            barrier()
            t5 = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev); dist.all_reduce(t5, op=dist.ReduceOp.MAX)
            r5 = torch.tensor([float(rays5)], dtype=torch.float64, device=dev); dist.all_reduce(r5, op=dist.ReduceOp.SUM)
            ok = None
            if rank == 0:
                film = grp.read_image(engine.FILM_SUM)
                ok = bool((film[:, 3] == c5spp).all() and np.isfinite(film).all())
                del film
            c5 = {"workload": f"C5: C2 scene, {c5w}x{c5h}, {c5spp} spp progressive ({per_gpu} per GPU; BASELINE's 4096 spp is the N = 8 case), passes of {k} spp per GPU, "
                              f"sample slices over {world} GPUs, film assembled on rank 0 (ncclReduce) at {checkpoints} spp",
                  "ms": float(t5[0]), "mrays_per_s": float(r5[0]) / float(t5[0]) / 1e3, "spp_per_s": c5spp / (float(t5[0]) / 1e3),
                  "spp_per_s_1080p_equiv": c5spp * (c5w * c5h) / (1920 * 1080) / (float(t5[0]) / 1e3), "assembly_ms_root_total": asm_ms,
                  "assemblies": len(checkpoints), "film_bytes": c5w * c5h * 16, "every_pixel_has_spp_samples": ok,
                  "rmse": "converged-image RMSE against the compiled reference is asserted by tests/test_golden.py::test_gpu_converged_image_rmse (4096 spp); "
                          "the 4K frame itself took the CPU reference 673 s for 8 rows (profiles/r01_c5_n8.json: 0.0280 vs floor 0.0281)"}
        except Exception as e:
            c5 = {"error": repr(e)}
        grp.set_scene(pos, nrm, uv)
        grp.set_camera(cam)

    # ---- e2e: host buffers through the C ABI, every step: scene upload + BVH build + render + image read-back ----
    ppos = torch.from_numpy(pos).pin_memory().numpy(); pnrm = torch.from_numpy(nrm).pin_memory().numpy(); puv = torch.from_numpy(uv).pin_memory().numpy()
    pimg = torch.empty((npix, 5), dtype=torch.float32).pin_memory().numpy()  # the caller's (pinned) image buffer, like Camera::mImage
    e2e_steps = max(1, min(args.steps, 3))
    e2e_warm = 1 if args.warmup > 0 else 0  # one untimed pass of the same loop body: first-use allocations (staging, read-back) are not a step
    e2e_rays = 0
    e2e_marks = [time.perf_counter()]  # host clock after every pass (this rank), the warm-up one included: reported as step_ms
    for it in range(e2e_warm + e2e_steps):
        if it == e2e_warm:
            barrier()
            t0 = time.perf_counter()
            e2e_rays = 0
            e2e_marks = [t0]
        if grp is not None:
            # the scene goes up once, on rank 0's PCIe link, and crosses NVLink to the other ranks; only rank 0 reads the image back
            root = rank == 0
            grp.set_scene(ppos if root else None, pnrm if root else None, puv if root else None, broadcast=True)
            grp.set_camera(cam)
            grp.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, split=engine.DIST_SPLIT_SAMPLES)
            grp.read_image(out=pimg)
        else:
            eng.set_scene(ppos, pnrm, puv)
            eng.set_camera(cam)
            eng.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)
            eng.read_image(out=pimg)
        e2e_rays += step_stats()["rays_traced"]
        e2e_marks.append(time.perf_counter())
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s, float(e2e_rays)], dtype=torch.float64, device="cuda")
    if world > 1:
        a = te.clone(); dist.all_reduce(a, op=dist.ReduceOp.MAX)
        b = te.clone(); dist.all_reduce(b, op=dist.ReduceOp.SUM)
        e2e_s, e2e_rays = float(a[0]), float(b[1])
    e2e = {"value": e2e_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(pos.nbytes + nrm.nbytes + uv.nbytes + 8 * 64 + 48),
           "d2h_bytes_per_step": int(pimg.nbytes), "steps": e2e_steps, "warmup": e2e_warm,
           "step_ms": [round((y - x) * 1e3, 3) for x, y in zip(e2e_marks, e2e_marks[1:])],
           "includes": "vmn_scene_set_triangles (H2D from pinned host) + vmn_scene_commit (GPU BVH build) + vmn_render + vmn_read_image (D2H)" if grp is None else
                       "vmn_dist_scene_set_triangles (H2D on rank 0 + ncclBroadcast) + vmn_dist_scene_commit (every GPU builds its BVH) + vmn_dist_render "
                       "(64 spp per rank + ncclReduce to rank 0) + vmn_dist_read_image (D2H on rank 0 only)"}
    build_ms = step_stats()["build_ms"]

    if rank != 0:
        grp.close()
        dist.destroy_process_group()
        return
    if grp is not None:  # rank 0 of an N > 1 run: the roofline of k_extend and the L2 probe use a plain context on this device
        eng = new_engine()
        eng.set_scene(pos, nrm, uv)

    # ---- roofline of the dominant kernel: measured visits per ray from a counting build on the same frame (untimed) ----
    ce = new_engine(counters=True)
    ce.set_scene(pos, nrm, uv)
    ce.set_camera(make_camera(width, height, 4))
    ce.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)
    cst = ce.stats()
    ce.close()
    nodes_per_ray = cst["nodes_visited"] / max(1, cst["rays_traced"])
    tris_per_ray = cst["tris_tested"] / max(1, cst["rays_traced"])
    b_ray = 32 + 8 + nodes_per_ray * cst["node_bytes"] + tris_per_ray * cst["tri_bytes"]
    peak, peak_src = peaks()
    achieved = (rays / max(1e-9, extend_ms / 1e3)) * b_ray / 1e9
    n_extend = args.steps * 5 * -(-(npix * args.spp) // (args.max_paths // npix * npix if args.max_paths >= npix else args.max_paths))
    # dram__bytes_read.sum + dram__bytes_write.sum per k_extend launch, from the committed ncu --set full capture (tools/update_traffic.py);
    # only valid for the kernel sources it was captured on: a stamp that does not match the sources of this build is refused
    traffic, traffic_note = None, "no capture"
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tr = json.load(f)["k_extend"]
        if tr.get("source_hash") == engine.kernel_source_hash():
            # captured on launches of tr["rays_per_launch"] rays; scaled to the rays one launch of THIS run processes (per launch, like `achieved`)
            traffic = tr["dram_bytes_per_launch"] / tr.get("rays_per_launch", 33177600) * (rays / max(1, n_extend))
            traffic_note = tr.get("source", "") + f" Captured at {tr.get('rays_per_launch')} rays per launch = {tr['dram_bytes_per_launch'] / tr.get('rays_per_launch', 33177600):.1f} B per ray; scaled to this run's rays per launch."
        else:
            traffic_note = f"stale capture refused: taken on sources {tr.get('source_hash')}, this build is {engine.kernel_source_hash()}"
    except Exception as e:
        traffic_note = f"no capture ({e!r})"
    roofline = {"bound": "hbm", "kernel": "k_extend", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": traffic_note, "peak_source": peak_src, "bytes_per_ray": b_ray, "nodes_per_ray": nodes_per_ray, "tris_per_ray": tris_per_ray,
                "node_bytes": cst["node_bytes"], "tri_bytes": cst["tri_bytes"], "extend_ms_per_step": extend_ms / args.steps,
                "extend_share_of_step": extend_ms / ms if world == 1 else None, "extend_launches": int(n_extend),
                "algorithmic_bytes_per_launch": b_ray * rays / max(1, n_extend),
                "note": "algorithmic bytes (SURVEY 8d B_ray); on C2 the ~60 MB of nodes+triangles stay L2-resident, so measured DRAM traffic is far below "
                        "the algorithmic bytes and the kernel is issue-bound (profiles/); tools/microbench_c3.py is the HBM-bound row"}

    # the scene is L2-resident on C2: the L2's read bandwidth, measured with the traversal's own access width over a 48 MiB working
    # set (vmn_probe_read_bandwidth), is the ceiling the algorithmic bytes actually run against; the same probe over 4 GiB reads HBM
    try:
        l2_gbs = eng.probe_read_bandwidth(48 << 20, 200)
        hbm_read_gbs = eng.probe_read_bandwidth(4 << 30, 2)
        roofline["l2"] = {"peak": l2_gbs, "unit": "GB/s", "achieved": achieved, "frac": achieved / l2_gbs, "hbm_read_probe": hbm_read_gbs,
                          "how": "vmn_probe_read_bandwidth: LDG.128 (.cg) reads, 148 x 8 CTAs x 256 threads, 8 loads in flight per thread; "
                                 "48 MiB working set x 200 passes (L2-resident), 4 GiB x 2 passes (HBM)"}
    except Exception as e:
        roofline["l2"] = {"peak": None, "error": repr(e)}

    # ---- C3, the traversal-roofline row of BASELINE.json (configs[2]): 16 M incoherent rays vs the 10 M-triangle soup — the scene that does
    # NOT fit the L2, so this is where the HBM roofline binds (C2 above is L2-resident). N = 1 only; same record as tools/microbench_c3.py.
    roofline_c3 = None
    if world == 1 and not args.no_c3:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import microbench_c3
            eng.close()  # free the C2 pool (3.6 GB) before the 10 M-triangle scene goes up
            eng = None
            roofline_c3 = microbench_c3.c3_record(stream, reps=5, check=0 if args.no_cpu else 200_000, device=local, flush=flush)
        except Exception as e:
            roofline_c3 = {"error": repr(e)}

    cpu = None
    if not args.no_cpu and world == 1:
        try:
            cpu = cpu_reference_sample(pos, nrm, uv, width, height, args.spp, budget_s=args.cpu_budget)
        except Exception as e:  # the checker is optional for the GPU number
            cpu = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "unavailable", "sample": repr(e)}

    out = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": WORKLOAD, "resolution": f"{width}x{height}", "spp_per_gpu": args.spp, "spp_total": spp_total,
                      "parallelism": f"sample-slice x{world} + ncclReduce of the film to rank 0 (vmn_dist)" if world > 1 else "single GPU",
                      "l2": "256 MiB L2 flush between timed steps", "paths_in_flight": args.max_paths},
           "spp_per_s": paths_all / npix / (ms / 1e3), "rays_per_step": rays_all / args.steps, "bvh_build_ms": build_ms,
           "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "roofline_c3": roofline_c3, "strong": strong,
           "c4": c4, "c5": c5, "cpu_baseline": cpu}
    if eng is not None:
        eng.close()
    if grp is not None:
        grp.close()
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(out) + "\n").encode())


if __name__ == "__main__":
    main()
