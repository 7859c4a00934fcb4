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


def reduce_film(film, group=None):
    """In-place all-reduce(SUM) of a film-sum tensor (any device; NCCL for CUDA tensors, gloo for CPU tensors)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(film, op=dist.ReduceOp.SUM, group=group)
    return film


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


def device_film_tensor(eng, device):
    """torch view (no copy) of the context's device film buffer, for NCCL."""
    import torch
    ptr, n = eng.film_device()

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}
    return torch.as_tensor(h, device=device)


def render_distributed(eng, device, mode, seed, split, rank, world, spp_total, max_segments=5, tile=64, group=None):
    """Render this rank's share on `eng`, all-reduce the film over the process group, return the RGBAZ image (every rank)."""
    eng.render(mode=mode, seed=seed, max_segments=max_segments, **render_params(split, rank, world, spp_total, tile))
    film = device_film_tensor(eng, device)
    reduce_film(film, group)
    return eng.resolve_device(film.data_ptr())
