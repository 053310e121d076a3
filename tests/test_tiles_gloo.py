"""CPU (-m "not gpu"): the N>1 path of bench.py / SURVEY §8e on the gloo backend, world_size 2 and 3.
Every rank fills only the pixels of its own screen tiles, packs them ([tile][th*tw][rgba,depth]), all ranks all_gather the
packed buffers, rank 0 unpacks the other ranks' tiles: the result must equal the frame rendered by one rank. The pack/unpack
used here are the numpy mirrors (package tiles.py) of pack_tiles_kernel; the -m gpu test compares the kernel with them."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tiles():
    spec = importlib.util.spec_from_file_location("hn_tiles", os.path.join(ROOT, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "tiles.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _full_frame(w, h):
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    frame = np.stack([x / w, y / h, (x + y) / (w + h), np.ones_like(x)], -1).astype(np.float32)
    depth = (1.0 + x * 0.01 + y * 0.02).astype(np.float32)
    return frame, depth


def _worker(rank, world, port, w, h, tw, th, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tiles = _tiles()
    full_f, full_d = _full_frame(w, h)
    own = tiles.owner_map(w, h, tw, th, world) == rank
    frame = np.where(own[..., None], full_f, 0).astype(np.float32)   # this rank "renders" only its tiles
    depth = np.where(own, full_d, 0).astype(np.float32)
    packed = torch.from_numpy(tiles.pack_tiles_host(frame, depth, tw, th, rank, world))
    assert packed.numel() == tiles.packed_floats(w, h, tw, th, world)
    gathered = [torch.zeros_like(packed) for _ in range(world)]
    dist.all_gather(gathered, packed)
    # max-over-ranks timing plumbing of bench.py
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = True
    if rank == 0:
        for r in range(1, world):
            tiles.unpack_tiles_host(gathered[r].numpy(), frame, depth, tw, th, r, world)
        ok = np.array_equal(frame, full_f) and np.array_equal(depth, full_d) and float(t) == world
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, bool(ok)))


@pytest.mark.parametrize("world,w,h", [(2, 200, 70), (3, 130, 100), (2, 64, 32)])
def test_tile_gather_gloo(world, w, h):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, w, h, 64, 32, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok in res)


def test_pack_roundtrip_single_process():
    tiles = _tiles()
    for world in (1, 2, 5, 8):
        w, h = 1920 // 8, 1080 // 8
        f, d = _full_frame(w, h)
        out_f, out_d = np.zeros_like(f), np.zeros_like(d)
        for r in range(world):
            tiles.unpack_tiles_host(tiles.pack_tiles_host(f, d, 64, 32, r, world), out_f, out_d, 64, 32, r, world)
        assert np.array_equal(out_f, f) and np.array_equal(out_d, d)


def _host_worker(rank, world, port, w, h, tw, th, q):
    """the sharded render-to-host path of bench.py (e2e, N > 1) on CPU: one shared host frame, every rank writes the tiles it owns"""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tiles = _tiles()
    names = [f"hn_test_{os.getpid()}_{port}" if rank == 0 else None]
    shared = tiles.SharedHostFrame(names[0], h, w, 4, create=True) if rank == 0 else None
    dist.broadcast_object_list(names, src=0)
    if rank != 0:
        shared = tiles.SharedHostFrame(names[0], h, w, 4, create=False)
    full_f, _ = _full_frame(w, h)
    own = tiles.owner_map(w, h, tw, th, world) == rank
    shared.array[own] = full_f[own]          # what scatter_tiles_kernel does with a host-mapped target
    dist.barrier()
    ok = np.array_equal(shared.array, full_f) if rank == 0 else True
    n_bytes_ok = shared.n_bytes == w * h * 16 and shared.host_ptr != 0
    dist.barrier()
    shared.close()
    dist.destroy_process_group()
    q.put((rank, bool(ok and n_bytes_ok)))


@pytest.mark.parametrize("world,w,h", [(2, 200, 70), (3, 130, 100)])
def test_shared_host_frame_gloo(world, w, h):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_host_worker, args=(r, world, port, w, h, 64, 32, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok in res)
