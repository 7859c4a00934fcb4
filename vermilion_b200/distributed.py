"""Multi-GPU host logic: one process per GPU, scene replicated, only the film (accumulation buffer) crosses NVLink.

The reference parallelises over pixels with OpenMP and no shared state but the disjoint film write
(core/integrators/pathtracer.cpp:226-227, 324). The two partitions of SURVEY.md §8e map that onto ranks:

  split="sample"  every rank renders the whole frame for its own slice of the sample sequence (FIXED mode only: the
                  reference's adaptive early-out is sequential per pixel). Film sums (sum r, sum g, sum b, n) add up, so
                  the exchange is ONE all-reduce(SUM) of width*height*4 floats.
  split="tile"    square tiles dealt round-robin (tile t -> rank t % world); a rank's film is zero outside its tiles, so the
                  same all-reduce(SUM) assembles the frame exactly (x + 0 = x) — works for COMPAT mode too.

There is no data-path collective inside the render itself. This module is pure host logic over torch.distributed and is
exercised on CPU with the gloo backend (tests/test_distributed_gloo.py); on the GPU box the same calls run over NCCL.
"""
import numpy as np


def sample_slice(rank, world, spp_total):
    """[begin, end) of the spp_total-long sample sequence for `rank`; slices differ by at most one sample."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world {world}")
    base, rem = divmod(spp_total, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def tile_owner_map(width, height, tile, world):
    """int32 [height, width]: rank owning each pixel under the interleaved tile split (same rule as vmn_render_params)."""
    tiles_x = -(-width // tile)
    ty, tx = np.meshgrid(np.arange(height) // tile, np.arange(width) // tile, indexing="ij")
    return ((ty * tiles_x + tx) % world).astype(np.int32)


def render_params(split, rank, world, spp_total, tile=64):
    """kwargs for Engine.render() describing this rank's share."""
    if split == "sample":
        b, e = sample_slice(rank, world, spp_total)
        return {"sample_begin": b, "sample_end": e}
    if split == "tile":
        return {"tile_size": tile, "tile_rank": rank, "tile_world": world}
    raise ValueError(f"unknown split {split!r}")


def reduce_film(film, group=None, force=False):
    """In-place all-reduce(SUM) of a film-sum tensor (any device; NCCL for CUDA tensors, gloo for CPU tensors).

    Stream order (CUDA): the collective is enqueued on torch's CURRENT stream, not on the engine's. Engine.render() returns with the
    engine's stream drained, so the film is complete when the collective starts; what follows the collective on the engine's stream
    (resolve, the next render's memset) must wait for it — use sync_after_reduce() or render_distributed(), which does. `force` runs
    the collective even in a world of one (tests)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and (force or dist.get_world_size(group) > 1):
        dist.all_reduce(film, op=dist.ReduceOp.SUM, group=group)
    return film


def sync_after_reduce(film):
    """Orders whatever the engine does next (on ITS stream) after a collective enqueued on torch's current stream: the host waits for
    that stream. An Engine created without stream= owns a non-blocking stream that nothing else orders against torch's."""
    if getattr(film, "is_cuda", False):
        import torch
        torch.cuda.current_stream(film.device).synchronize()


def resolve(film_sum):
    """Film sums -> RGBAZ rows exactly as k_resolve / pathtracer.cpp:318-323 do (numpy, host side; tests and tools only)."""
    f = np.asarray(film_sum, np.float32).reshape(-1, 4)
    out = np.zeros((len(f), 5), np.float32)
    n = f[:, 3]
    ok = n > 0
    with np.errstate(invalid="ignore", divide="ignore"):
        rgb = f[:, :3] / n[:, None]
    out[ok, :3] = np.clip(rgb[ok], 0.0, 1.0)
    out[ok, 3] = 1.0
    out[ok, 4] = n[ok]
    return out


def resolve_bruteforce(film_sum, film_aux):
    """BRUTEFORCE film planes -> RGBAZ rows as k_resolve_bf / integrators.cpp:176-183 do: std::max(std::min(x, 1), 0) keeps a NaN mean,
    alpha = sum alpha / n, depth = distance of the last primary RayCast (numpy, host side; tests and tools only)."""
    f = np.asarray(film_sum, np.float32).reshape(-1, 4)
    a = np.asarray(film_aux, np.float32).reshape(-1, 2)
    out = np.zeros((len(f), 5), np.float32)
    n = f[:, 3]
    ok = n > 0
    with np.errstate(invalid="ignore", divide="ignore"):
        rgb = f[:, :3] / n[:, None]
        m = np.where(1.0 < rgb, np.float32(1.0), rgb)      # std::min(x, 1.f) = (1 < x) ? 1 : x
        rgb = np.where(m < 0.0, np.float32(0.0), m)         # std::max(m, 0.f) = (m < 0) ? 0 : m
        alpha = a[:, 0] / n
    out[ok, :3] = rgb[ok]
    out[ok, 3] = alpha[ok]
    out[ok, 4] = a[ok, 1]
    return out


def device_film_tensor(eng, device):
    """torch view (no copy) of the context's device film buffer, for NCCL. The view is only valid until the next Engine.set_camera
    with a different resolution (vmn_camera_set reallocates the film): take a new view after changing the camera."""
    import torch
    ptr, n = eng.film_device()

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}
    return torch.as_tensor(h, device=device)


def device_film_aux_tensor(eng, device):
    """torch view (no copy) of the second plane of a BRUTEFORCE film (sum alpha, last distance), for NCCL."""
    import torch
    ptr, n = eng.film_aux_device()

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}
    return torch.as_tensor(h, device=device)


def render_bruteforce_distributed(eng, device, seed, rank, world, tile=64, shadow=False, group=None, force_reduce=False):
    """BruteForceTracer across ranks: interleaved tiles (its early-out is sequential per pixel, so pixels — not samples — are split);
    both film planes are zero outside a rank's tiles, so two all-reduce(SUM) assemble the frame exactly. Returns RGBAZ (every rank)."""
    from . import engine as _engine
    eng.render(mode=_engine.MODE_BRUTEFORCE, seed=seed, shadow=shadow, **render_params("tile", rank, world, 0, tile))
    film = device_film_tensor(eng, device)
    aux = device_film_aux_tensor(eng, device)
    reduce_film(film, group, force_reduce)
    reduce_film(aux, group, force_reduce)
    sync_after_reduce(aux)  # the resolve below runs on the engine's stream, the collectives ran on torch's
    return eng.resolve_device(film.data_ptr())  # the context resolves with its own (now reduced) second plane


def render_distributed(eng, device, mode, seed, split, rank, world, spp_total, max_segments=5, tile=64, group=None, force_reduce=False):
    """Render this rank's share on `eng`, all-reduce the film over the process group, return the RGBAZ image (every rank). Safe for an
    Engine on any stream: the render drains the engine's stream before the collective, and the host waits for the collective's stream
    before the resolve (and the next render's film memset) is enqueued on the engine's."""
    eng.render(mode=mode, seed=seed, max_segments=max_segments, **render_params(split, rank, world, spp_total, tile))
    film = device_film_tensor(eng, device)
    reduce_film(film, group, force_reduce)
    sync_after_reduce(film)
    return eng.resolve_device(film.data_ptr())
