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


class SharedHostFrame:
    """One frame in host memory that every rank's process of a box has mapped (POSIX shared memory): the target of the sharded
    render-to-host path - each rank page-locks the mapping (Testbed.host_register) and scatters the tiles it owns into it over its own
    PCIe link (Testbed.scatter_tiles); any rank can read `array` once all ranks have synchronised their streams and passed a barrier.

    create=True on one rank (it owns and finally unlinks the segment), create=False with the same name on the others."""

    def __init__(self, name: str, height: int, width: int, channels: int = 4, create: bool = False):
        from multiprocessing import shared_memory
        n_bytes = height * width * channels * 4
        self._shm = shared_memory.SharedMemory(name=name, create=create, size=n_bytes if create else 0)
        self._owner = create
        if not create:
            # Python < 3.13 registers attached segments with the resource tracker too, which then unlinks them (and warns) when this
            # process exits although the creator still owns them
            try:
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self._shm._name, "shared_memory")
            except Exception:
                pass
        self.name = name
        self.n_bytes = n_bytes
        self.array = np.ndarray((height, width, channels), np.float32, buffer=self._shm.buf)
        if create:
            self.array[...] = 0.0

    @property
    def host_ptr(self) -> int:
        return self.array.ctypes.data

    def close(self):
        self.array = None
        try:
            self._shm.close()
        except BufferError:
            pass
        if self._owner:
            try:
                self._shm.unlink()
            except FileNotFoundError:
                pass
