"""Screen-tile sharding of one frame over the GPUs of a box: host-side layout arithmetic only.

Tile t (row-major over ceil(W/tw) x ceil(H/th) tiles) belongs to rank t % world — the same rule the kernels apply
(csrc/hn_march.cu tile_owned, csrc/hn_mesh.cu mesh_tile_owned, csrc/hn_composite.cu pack_tiles_kernel). The packed
exchange buffer of a rank is [local tile][th*tw pixels][5 floats = rgba, depth]; rank r's local tile k is global tile
k*world + r. The reference's analogue is the whole-frame cudaMemcpyPeerAsync in Testbed::use_device
(src/testbed.cu:5604-5605); there is no intra-frame tiling in the reference.

The numpy pack/unpack below mirror the device kernels for the CPU (gloo) tests of the multi-rank path; the GPU path
never calls them.
"""
from __future__ import annotations

import numpy as np

FLOATS_PER_PIXEL = 5


def tile_grid(width: int, height: int, tile_w: int, tile_h: int) -> tuple[int, int]:
    return (width + tile_w - 1) // tile_w, (height + tile_h - 1) // tile_h


def tile_count(width: int, height: int, tile_w: int, tile_h: int, rank: int, world: int) -> int:
    tx, ty = tile_grid(width, height, tile_w, tile_h)
    return (tx * ty + world - 1 - rank) // world


def packed_floats(width: int, height: int, tile_w: int, tile_h: int, world: int) -> int:
    """Size (in floats) of ONE rank's exchange buffer; every rank uses rank 0's (largest) size so all_gather is uniform."""
    return tile_count(width, height, tile_w, tile_h, 0, world) * tile_w * tile_h * FLOATS_PER_PIXEL


def owner_map(width: int, height: int, tile_w: int, tile_h: int, world: int) -> np.ndarray:
    """int32 [H, W]: owning rank of every pixel."""
    tx, _ = tile_grid(width, height, tile_w, tile_h)
    y, x = np.mgrid[0:height, 0:width]
    return (((y // tile_h) * tx + (x // tile_w)) % world).astype(np.int32)


def pack_tiles_host(frame: np.ndarray, depth: np.ndarray, tile_w: int, tile_h: int, rank: int, world: int) -> np.ndarray:
    h, w = depth.shape
    tx, ty = tile_grid(w, h, tile_w, tile_h)
    out = np.zeros(packed_floats(w, h, tile_w, tile_h, world), np.float32).reshape(-1, tile_h, tile_w, FLOATS_PER_PIXEL)
    for k in range(tile_count(w, h, tile_w, tile_h, rank, world)):
        t = k * world + rank
        x0, y0 = (t % tx) * tile_w, (t // tx) * tile_h
        x1, y1 = min(x0 + tile_w, w), min(y0 + tile_h, h)
        out[k, : y1 - y0, : x1 - x0, :4] = frame[y0:y1, x0:x1]
        out[k, : y1 - y0, : x1 - x0, 4] = depth[y0:y1, x0:x1]
    return out.reshape(-1)


def unpack_tiles_host(packed: np.ndarray, frame: np.ndarray, depth: np.ndarray, tile_w: int, tile_h: int, rank: int, world: int) -> None:
    h, w = depth.shape
    tx, ty = tile_grid(w, h, tile_w, tile_h)
    p = packed.reshape(-1, tile_h, tile_w, FLOATS_PER_PIXEL)
    for k in range(tile_count(w, h, tile_w, tile_h, rank, world)):
        t = k * world + rank
        x0, y0 = (t % tx) * tile_w, (t // tx) * tile_h
        x1, y1 = min(x0 + tile_w, w), min(y0 + tile_h, h)
        frame[y0:y1, x0:x1] = p[k, : y1 - y0, : x1 - x0, :4]
        depth[y0:y1, x0:x1] = p[k, : y1 - y0, : x1 - x0, 4]


# ---- data-parallel training (SURVEY §8e "Training"): the ray stream of one step is cut into `world` consecutive blocks ----
def train_ray_shard(n_rays_per_rank: int, rank: int, world: int, n_images: int) -> tuple[np.ndarray, np.ndarray]:
    """Global ray indices of this rank's block and the image each ray samples — the rule of train_raygen_kernel
    (csrc/hn_train.cu): global ray g = rank * n + i, image = (g * n_images) // (n * world) % n_images. The loss of every ray is
    normalised by the GLOBAL ray count, so the SUM all-reduce of the per-rank gradient buffers is the gradient of the whole
    batch and hn_train_apply takes grad_scale = 1."""
    g = rank * n_rays_per_rank + np.arange(n_rays_per_rank, dtype=np.int64)
    img = (g * n_images) // (n_rays_per_rank * world) % n_images
    return g, img.astype(np.int64)
