"""CPU, world_size 2 over gloo: the tile partition and the film reduce of the multi-GPU path
(mitsuba_b200/dist.py).  Each rank splats the samples of ITS tiles into a bordered film (with the
oracle's ImageBlock::put, here only a stand-in for a rank's renderer); the reduced film must equal
the film of a single process that splats every sample -- including the filter borders that
reach into tiles owned by the other rank (imageblock.h:103-107)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mitsuba_b200 import dist as mdist
from mitsuba_b200 import scenes


def _samples(w, h, spp, seed=3):
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:h, 0:w]
    px = np.repeat(xs.ravel(), spp); py = np.repeat(ys.ravel(), spp)
    s = np.zeros((px.size, 6), np.float32)
    s[:, 0] = px + rng.random(px.size); s[:, 1] = py + rng.random(px.size)
    s[:, 2:5] = rng.random((px.size, 3)); s[:, 5] = 1.0
    return s, px, py


def _worker(rank, world, port, w, h, spp, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    orc = oracle.Oracle(scenes.cbox(width=w, height=h, spp=spp))
    s, px, py = _samples(w, h, spp)
    tx = (w + 31) // 32
    mine = np.isin((py // 32) * tx + (px // 32), mdist.tiles_of_rank(w, h, rank, world))
    orc.film_clear()
    orc.film_put(s[mine])
    film = torch.from_numpy(orc.film_download().copy())
    mdist.reduce_film(film, dst=0)
    if rank == 0:
        np.save(out, film.numpy())
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_partition_covers_every_tile_once():
    for (w, h) in ((1920, 1080), (512, 512), (50, 37), (33, 1)):
        for world in (1, 2, 3, 4, 8):
            tiles = sorted(t for r in range(world) for t in mdist.tiles_of_rank(w, h, r, world))
            assert tiles == list(range(((w + 31) // 32) * ((h + 31) // 32)))
            assert sum(mdist.pixels_of_rank(w, h, r, world) for r in range(world)) == w * h
    for spp in (1, 7, 256):
        for world in (1, 2, 3, 8):
            r = [mdist.sample_range_of_rank(spp, k, world) for k in range(world)]
            assert r[0][0] == 0 and sum(c for _, c in r) == spp
            assert all(r[k][0] + r[k][1] == r[k + 1][0] for k in range(world - 1))


@pytest.mark.timeout(180)
def test_two_rank_film_reduce_equals_single_process(tmp_path, oracle_mod):
    w, h, spp = 80, 48, 2
    out = str(tmp_path / "film.npy")
    mp.spawn(_worker, args=(2, _free_port(), w, h, spp, out), nprocs=2, join=True)
    reduced = np.load(out)
    orc = oracle_mod.Oracle(scenes.cbox(width=w, height=h, spp=spp))
    s, _, _ = _samples(w, h, spp)
    orc.film_clear(); orc.film_put(s)
    whole = orc.film_download()
    assert reduced.shape == whole.shape
    assert np.allclose(reduced, whole, rtol=1e-5, atol=1e-6)
    assert np.allclose(mdist.develop(reduced, 2), orc.film_develop(), rtol=1e-5, atol=1e-6)
