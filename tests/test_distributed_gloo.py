"""CPU tests of the multi-GPU host logic (vermilion_b200/distributed.py) with world_size 2 over gloo. The per-rank film
sums come from the CPU oracle (the checker), so what is tested is the partition + exchange + resolve logic the GPU path uses
unchanged over NCCL: sample-slice and tile splits must reassemble the single-process frame."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import checker
from vermilion_b200 import distributed as vd
from vermilion_b200 import scenes

W, H, SPP, SEED = 40, 24, 12, 5


def _cam():
    cs = checker.camera_settings(W, H, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 2.16), SPP)
    return checker.camera_fields_py(cs)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, split, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pos, nrm, uv = scenes.spheres(2, 1, 2, 150.0)
        orc = checker.Oracle(pos, nrm, uv)
        cam = _cam()
        if split == "sample":
            p = vd.render_params("sample", rank, world, SPP)
            _, raw, _ = orc.render(cam, checker.MODE_FIXED, checker.RNG_PHILOX, SEED, s0=p["sample_begin"], s1=p["sample_end"], want_raw=True)
        else:
            _, raw, _ = orc.render(cam, checker.MODE_FIXED, checker.RNG_PHILOX, SEED, want_raw=True)
            p = vd.render_params("tile", rank, world, SPP, tile=8)
            own = vd.tile_owner_map(W, H, p["tile_size"], p["tile_world"]).reshape(-1) == p["tile_rank"]
            raw[~own] = 0.0
        film = torch.from_numpy(raw.reshape(-1).copy())
        vd.reduce_film(film)
        np.save(os.path.join(out_dir, f"{split}_{rank}.npy"), film.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("split", ["sample", "tile"])
def test_two_ranks_reassemble_the_frame(tmp_path, split):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), split, str(tmp_path)), nprocs=world, join=True)
    films = [np.load(tmp_path / f"{split}_{r}.npy").reshape(-1, 4) for r in range(world)]
    assert np.array_equal(films[0], films[1])  # every rank holds the reduced film
    pos, nrm, uv = scenes.spheres(2, 1, 2, 150.0)
    img, raw, _ = checker.Oracle(pos, nrm, uv).render(_cam(), checker.MODE_FIXED, checker.RNG_PHILOX, SEED, want_raw=True)
    assert np.array_equal(films[0][:, 3], raw[:, 3])  # sample counts
    if split == "tile":
        assert np.array_equal(films[0], raw)  # x + 0 = x: exact
        assert np.array_equal(vd.resolve(films[0]), img)
    else:  # fp32 sums in a different association order
        assert np.allclose(films[0], raw, rtol=1e-6, atol=1e-6)
        assert np.abs(vd.resolve(films[0]) - img).max() < 1e-6
    assert raw[:, :3].max() > 0


def _worker_bf(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pos, nrm, uv = scenes.cornell()
        cs = checker.camera_settings(W, H, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 2.16), SPP)
        _, _, _, raw = checker.Oracle(pos, nrm, uv, spheres=False).render_bruteforce(checker.camera_fields_py(cs), seed=SEED, want_raw=True)
        p = vd.render_params("tile", rank, world, 0, tile=8)
        own = vd.tile_owner_map(W, H, p["tile_size"], p["tile_world"]).reshape(-1) == p["tile_rank"]
        raw[~own] = 0.0  # what a rank's two film planes hold outside its tiles
        film, aux = torch.from_numpy(raw[:, :4].copy()), torch.from_numpy(raw[:, 4:].copy())
        vd.reduce_film(film)
        vd.reduce_film(aux)
        np.save(os.path.join(out_dir, f"bf_{rank}.npy"), vd.resolve_bruteforce(film.numpy(), aux.numpy()))
    finally:
        dist.destroy_process_group()


def test_two_ranks_reassemble_the_bruteforce_frame(tmp_path):
    """BruteForceTracer splits by tiles; its second film plane (sum alpha, last distance — inf where the open Cornell box lets a primary
    ray escape) is reduced with the first; the resolved frame equals the single-process oracle image bit for bit."""
    world = 2
    mp.spawn(_worker_bf, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    imgs = [np.load(tmp_path / f"bf_{r}.npy") for r in range(world)]
    pos, nrm, uv = scenes.cornell()
    cs = checker.camera_settings(W, H, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 2.16), SPP)
    want, _, _ = checker.Oracle(pos, nrm, uv, spheres=False).render_bruteforce(checker.camera_fields_py(cs), seed=SEED)
    for im in imgs:
        assert np.array_equal(im.view(np.uint32), want.view(np.uint32))
    assert (want[:, 3] > 0).any() and want[:, :3].max() > 0


def test_partitions_cover_exactly_once():
    for world in (1, 2, 3, 4, 8):
        for spp in (1, 7, 64, 4096):
            sl = [vd.sample_slice(r, world, spp) for r in range(world)]
            assert sl[0][0] == 0 and sl[-1][1] == spp and all(sl[i][1] == sl[i + 1][0] for i in range(world - 1))
            assert max(e - b for b, e in sl) - min(e - b for b, e in sl) <= 1
        own = vd.tile_owner_map(100, 37, 16, world)
        assert own.min() == 0 and own.max() == min(world - 1, own.max()) and set(np.unique(own)) <= set(range(world))
        assert own[0, 0] == 0 and own[0, 16] == 1 % world and own[16, 0] == 7 % world  # tiles_x = 7
    with pytest.raises(ValueError):
        vd.sample_slice(2, 2, 8)
    with pytest.raises(ValueError):
        vd.render_params("rows", 0, 1, 8)
