"""CPU (-m "not gpu"): the N>1 path of the training step on the gloo backend, world_size 2.
Each rank takes its block of the step's ray stream (tiles.train_ray_shard = the rule of train_raygen_kernel), computes the
gradients of ITS rays normalised by the global ray count (here with the CPU oracle standing in for the CUDA kernels), the
ranks SUM all-reduce the fp32 gradient buffers: the result must equal the one-rank gradient of the whole batch."""
import ctypes as C
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
f32 = np.float32


def _mod(name, file):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "surface-irradiance-estimation-from-neural-radiance-fields_b200", file))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _batch(n_total):
    rng = np.random.default_rng(17)
    o = rng.normal(size=(n_total, 3)).astype(f32)
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * f32(1.2) + f32(0.5)
    tgt = f32(0.5) + (rng.random((n_total, 3), dtype=f32) - f32(0.5)) * f32(0.5)
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    t0 = np.full(n_total, 0.2, f32)
    return o, d, t0, rng.random((n_total, 3), dtype=f32), rng.random((n_total, 3), dtype=f32)


def _grads(orc, synthetic, idx, n_total):
    cfg = orc.net_config(8, 4, 19, 16, 2.0, 1, 2)

    class Info:
        n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = 8, 4, 1, 2
        n_params = orc.lib.orc_n_params(C.byref(cfg))
        grid_param_offset = orc.lib.orc_grid_param_offset(C.byref(cfg))

    params = synthetic.random_params(Info, 1337, 0.5)
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    p = orc.render_params(8, 8, synthetic.default_camera())
    o, d, t0, tg, bg = _batch(n_total)
    # the oracle normalises by the rays it is given: rescale to the global count through loss_scale
    scale = 128.0 * len(idx) / n_total
    loss, g, _ = orc.train_gradients(p, cfg, params, bits, o[idx], d[idx], t0[idx], tg[idx], bg[idx], loss_type=1, loss_scale=scale)
    return loss * len(idx) / n_total, g


def _worker(rank, world, port, n_per_rank, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import orc
    tiles, synthetic = _mod("hn_tiles", "tiles.py"), _mod("hn_synthetic", "synthetic.py")
    g_idx, img = tiles.train_ray_shard(n_per_rank, rank, world, 5)
    loss, grads = _grads(orc, synthetic, g_idx, n_per_rank * world)
    buf = torch.from_numpy(grads)
    dist.all_reduce(buf)                      # what bench.py --workload config5 does with NCCL on the device buffer
    lt = torch.tensor([loss], dtype=torch.float64)
    dist.all_reduce(lt)
    if rank == 0:
        q.put((buf.numpy().copy(), float(lt.item()), g_idx.tolist(), img.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_ray_shards_partition_the_stream():
    tiles = _mod("hn_tiles", "tiles.py")
    for world in (1, 2, 3, 8):
        n = 128
        all_g, all_img = [], []
        for r in range(world):
            g, img = tiles.train_ray_shard(n, r, world, 7)
            all_g += g.tolist(); all_img += img.tolist()
        assert all_g == list(range(n * world))                       # disjoint, complete, ordered
        one_g, one_img = tiles.train_ray_shard(n * world, 0, 1, 7)
        assert all_img == one_img.tolist()                           # same image per global ray as a single rank would pick
        assert set(all_img) == set(range(7))


def test_gradient_allreduce_equals_full_batch_world2():
    world, n_per_rank = 2, 6
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29650 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_per_rank, q)) for r in range(world)]
    for p in procs:
        p.start()
    summed, loss, g_idx, _ = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    import orc
    synthetic = _mod("hn_synthetic", "synthetic.py")
    full_loss, full = _grads(orc, synthetic, np.arange(world * n_per_rank), world * n_per_rank)
    assert g_idx == list(range(n_per_rank))
    assert abs(loss - full_loss) <= 1e-5 * abs(full_loss)
    assert np.abs(summed - full).max() <= 1e-5 * np.abs(full).max()
