"""Multi-GPU partition of a render: one process per GPU, tiles interleaved over ranks, one
reduce of the film at the end.

The reference merges blocks into the film with `ImageBlock::put(block)` -- an overlap-add that
includes the filter border (include/mitsuba/render/imageblock.h:103-107, called from
src/librender/renderproc.cpp:149-156).  Across GPUs the same merge is a SUM of the bordered
per-rank films: `torch.distributed.reduce(film, dst=0, op=SUM)` (NCCL over NVLink on the GPU
box, gloo in the CPU tests).  There is no other communication on this path.
"""
import numpy as np

TILE = 32


def tiles_of_rank(width, height, rank, world):
    """Tile ids (row-major over ceil(w/32) x ceil(h/32)) rendered by `rank`: t % world == rank
    -- the rule mtsb_render applies for (tile_index, tile_count)."""
    tx, ty = (width + TILE - 1) // TILE, (height + TILE - 1) // TILE
    return [t for t in range(tx * ty) if t % world == rank]


def pixels_of_rank(width, height, rank, world):
    tx = (width + TILE - 1) // TILE
    n = 0
    for t in tiles_of_rank(width, height, rank, world):
        ox, oy = (t % tx) * TILE, (t // tx) * TILE
        n += min(TILE, width - ox) * min(TILE, height - oy)
    return n


def sample_range_of_rank(spp, rank, world):
    """Sample-split alternative for images with few tiles: rank renders samples [first, first+count)
    (passed to mtsb_render as spp_offset / spp); the counter-based RNG is keyed by the global
    sample index, so the union over ranks equals a single-GPU render."""
    base, rem = divmod(spp, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


class _DevPtr:
    """Zero-copy torch view of device memory owned by the library (the film)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def film_tensor(scene):
    import torch
    ptr, n = scene.film_device_ptr()
    return torch.as_tensor(_DevPtr(ptr, n), device="cuda")


def reduce_film(film, dst=0):
    """Sum the per-rank bordered films onto `dst` (in place).  `film` is a torch tensor: the
    device film (NCCL) or a host copy (gloo)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def develop(film5, border):
    """HDRFilm::develop on a host copy of the bordered film: rgb = sum(w L) / sum(w)."""
    f = np.asarray(film5)
    h, w = f.shape[0] - 2 * border, f.shape[1] - 2 * border
    c = f[border:border + h, border:border + w]
    wgt = c[..., 4:5]
    return np.where(wgt != 0, c[..., :3] / np.where(wgt != 0, wgt, 1), 0).astype(np.float32)
