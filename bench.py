#!/usr/bin/env python
"""Headline benchmark: Mrays/s (and ms/frame) of one 1920x1080 hybrid NeRF + 1 Mi-triangle-mesh frame with sun shadow rays
(BASELINE.json configs[2]) on N B200s, frame tile-sharded across ranks and gathered over NCCL.

  python bench.py [--gpus N] [--steps K] [--warmup W]           our CUDA path (one process per GPU under torchrun for N > 1)
  python bench.py --impl reference ...                          the reference's path on the host cores (CPU oracle)

Prints ONE JSON line (rank 0). A step = one frame. `value` is measured with scene state resident in HBM and the frame left in
device memory; `e2e` goes through the pyngp-style Testbed.render() call with host buffers (camera H2D, RGBA frame D2H).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, H = 1920, 1080
MESH_NU, MESH_NV = 1024, 512  # 1024 * 512 * 2 = 1 048 576 triangles
ALGO_BYTES_PER_SAMPLE = 512    # L*8 corners*F*2 B gathered per network sample
ALGO_FLOP_PER_SAMPLE = 18816
NCU_EVAL_DRAM_BYTES = 186.7e6  # dram__bytes_read.sum + dram__bytes_write.sum of one eval_tc_kernel launch (ncu --set full, profiles/r01_ncu_eval_tc_v5.txt)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0))), "measured"
    return 6650.0, 1400.0, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.idx = gpu_index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower() == "active":
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=3)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def build_scene_inputs(hn):
    mesh = hn.synthetic.torus_mesh(MESH_NU, MESH_NV)
    grid = hn.synthetic.sphere_density_grid(0.35)
    cam = hn.synthetic.default_camera()
    return mesh, grid, cam


def reference_arm(args):
    """--impl reference: the reference has no CPU implementation of this path and its CUDA build cannot be produced here
    (DESIGN.md); the timed baseline is the scalar host oracle (restatement of the reference pipeline, oracle/) on all host
    cores. Each step renders a bounded sample: the same camera/scene at 1/8 x 1/8 resolution (1/64 of the rays)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import importlib
    import orc
    synthetic = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200.synthetic") if False else None
    # the oracle leg must not need the CUDA library: load synthetic.py directly
    import importlib.util
    spec = importlib.util.spec_from_file_location("hn_synthetic", os.path.join(ROOT, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "synthetic.py"))
    synthetic = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synthetic)

    cores = os.cpu_count() or 1
    w, h = W // 8, H // 8
    cfg = orc.net_config(8, 4, 19, 16, float(np.exp(np.float32(np.log(np.float32(128.0)) / np.float32(7.0)))), 1, 2)

    class Info:  # shape of hn_network_info for synthetic.random_params
        n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = 8, 4, 1, 2
        n_params = orc.lib.orc_n_params(orc.C.byref(cfg))
        grid_param_offset = orc.lib.orc_grid_param_offset(orc.C.byref(cfg))

    params = synthetic.random_params(Info, 1337, 1e-4)
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    mesh = synthetic.torus_mesh(MESH_NU, MESH_NV)
    scene = orc.Scene([mesh])
    bb = scene.aabb()
    aabb = np.concatenate([bb[:3] - np.float32(4.0), bb[3:] + np.float32(4.0)])
    # bound the whole run (warm-up + steps) to about 150 s of CPU work whatever K and W are: one 96x54 calibration frame gives
    # the host's ray rate, the per-step sample is the same view at the 16:9 resolution that fits the budget (at most 240x135)
    pc = orc.render_params(96, 54, synthetic.default_camera(), mesh_aabb=aabb, n_threads=cores, snap=True)
    t0 = time.perf_counter()
    orc.render_frame(pc, cfg, params, bits, scene)
    rate = 96 * 54 / (time.perf_counter() - t0)
    rays = min(w * h, max(64 * 36, rate * 150.0 / max(1, args.warmup + args.steps)))
    h = max(36, int((rays * 9 / 16) ** 0.5))
    w = h * 16 // 9
    p = orc.render_params(w, h, synthetic.default_camera(), mesh_aabb=aabb, n_threads=cores, snap=True)
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        orc.render_frame(p, cfg, params, bits, scene)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    mrays = (w * h) / (ms * 1e-3) / 1e6
    line = {
        "impl": "reference", "metric": "Mrays/s, 1080p hybrid NeRF+mesh frame (primary rays)", "value": mrays, "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "fp16 params/activations, fp32 accumulate", "data": "synthetic",
        "config": {"workload": "configs[2]: 1920x1080 hybrid NeRF (base.json, 128^3 sphere occupancy) + 1 048 576-triangle torus, sun shadow rays",
                   "sample": f"{w}x{h} frame of the same view per step ({w * h / (W * H):.5f} of the rays; sized for ~150 s of CPU work in total)"},
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": f"{w}x{h} frame ({w * h / (W * H):.5f} of the 1080p rays), {args.steps} steps"},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


class _DevArray:
    """Raw device pointer -> torch tensor (torch.as_tensor reads __cuda_array_interface__); no copy."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def train_arm(args):
    """--workload config5 (BASELINE.json configs[4]): NeRF training steps at target batch 2^18 samples per GPU on a synthetic
    in-memory dataset; data-parallel over rays, fp32 gradient buffer summed with one NCCL all-reduce per step (weak scaling)."""
    import importlib
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
    tb = hn.Testbed(device=local)
    tb.reload_network_from_file(os.path.join(ROOT, "configs", "nerf", "base.json"))
    info = tb.network_info()
    tb.set_params(hn.synthetic.random_params(info, 1337, 1e-4))
    tb.set_density_grid(hn.synthetic.sphere_density_grid(0.35), 1)
    imgs, cams, focal = hn.synthetic.sphere_dataset(64, 256)
    tb.set_training_data(imgs, cams, focal)
    tb.set_tile_shard(rank, world)
    tb.reset_trainer(1337)
    ptr, n = tb.train_gradient_buffer()
    grads = torch.as_tensor(_DevArray(ptr, n), device=dev)

    def step():
        loss = tb.train_accumulate(0)
        if world > 1:
            torch.cuda.synchronize()  # the library works on its own stream
            dist.all_reduce(grads)
            torch.cuda.synchronize()
        tb.train_apply(1.0)
        return loss

    for _ in range(max(args.warmup, 8)):  # lets the rays-per-batch controller settle on the 2^18-sample target
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    samples = 0
    e0.record()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        loss = step()
        samples += tb.train_stats().n_samples_last
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if world > 1:
        t = torch.tensor([ms, float(samples)], device=dev, dtype=torch.float64)
        tm = t.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ts = t.clone(); dist.all_reduce(ts)
        ms, samples = float(tm[0].item()), float(ts[1].item())
    clocks = sampler.stop() if rank == 0 else None
    # where a multi-GPU step goes (host clock around host-synchronous phases, max over ranks): compute, wait for the slowest rank,
    # the all-reduce itself (entered together, right after a barrier), optimizer
    phases = None
    if world > 1:
        acc = [0.0, 0.0, 0.0, 0.0]
        n_ph = 10
        for _ in range(n_ph):
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter(); tb.train_accumulate(0); torch.cuda.synchronize()
            t1 = time.perf_counter(); dist.barrier(); torch.cuda.synchronize()
            t2 = time.perf_counter(); dist.all_reduce(grads); torch.cuda.synchronize()
            t3 = time.perf_counter(); tb.train_apply(1.0); torch.cuda.synchronize()
            t4 = time.perf_counter()
            for i, v in enumerate((t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
                acc[i] += v * 1e3 / n_ph
        tp = torch.tensor(acc, device=dev, dtype=torch.float64)
        dist.all_reduce(tp, op=dist.ReduceOp.MAX)
        phases = {k: round(float(v), 4) for k, v in zip(("accumulate", "wait_slowest_rank", "all_reduce", "apply"), tp.tolist())}
    tb.set_option("profile", 1.0)
    step()
    st = tb.train_stats()
    if rank == 0:
        line = {"metric": "Msamples/s, NeRF training step (forward + backward + Adam), target batch 2^18 samples per GPU", "value": samples / args.steps / (ms * 1e-3) / 1e6,
                "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 8), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "fp16 params/activations, fp32 accumulate + fp32 master weights", "data": "synthetic",
                "config": {"workload": "configs[4]: train_nerf_step, configs/nerf/base.json, 64 analytic views 256x256, Huber loss, Adam/ExponentialDecay/EMA, "
                                       "gradient all-reduce (fp32, 46.8 MB) over NCCL", "rays_per_batch_per_gpu": int(st.rays_per_batch),
                           "samples_per_step_all_ranks": samples / args.steps, "l2": "per-step working set (activations of 2^18 samples, 47 MB gradients) exceeds reuse across steps"},
                "clocks": clocks, "gpu_launches": int(args.steps * 45), "loss": loss,
                "e2e": {"value": samples / args.steps / (wall_ms * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": wall_ms, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 8},
                "kernel_ms": {"sample": st.ms_sample, "forward": st.ms_forward, "loss": st.ms_loss, "backward": st.ms_backward, "optimizer": st.ms_optimizer},
                "step_phases_ms": phases, "roofline": None, "cpu_baseline": None}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--eval-impl", type=int, default=1, help="network kernel: 1 = tcgen05/TMEM (default), 0 = warp MMA")
    ap.add_argument("--clip", action="store_true", help="stop marching at the mesh depth instead of the reference's per-pixel test")
    ap.add_argument("--workload", default="config3", choices=["config2", "config3", "config4", "config5"],
                    help="BASELINE.json configs[1] NeRF-only 1080p / configs[2] hybrid 1080p (default, the headline) / configs[3] 4K hybrid + envmap light")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    global W, H
    if args.workload == "config4":
        W, H = 3840, 2160
    if args.impl == "reference":
        return reference_arm(args)
    if args.workload == "config5":
        return train_arm(args)

    import importlib
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
    mesh, grid, cam = build_scene_inputs(hn)
    tb = hn.Testbed(device=local)
    tb.reload_network_from_file(os.path.join(ROOT, "configs", "nerf", "base.json"))
    info = tb.network_info()
    tb.set_params(hn.synthetic.random_params(info, 1337, 1e-4))
    tb.set_density_grid(grid, 1)
    if args.workload != "config2":
        tb.add_mesh(mesh)
        tb.build_scene()
    if args.workload == "config4":
        tb.set_envmap(hn.synthetic.envmap(1024, 512, seed=3))
        tb.set_option("env_light_samples", 16.0)
    tb.snap_to_pixel_centers = True
    tb.set_option("depth_merge_clip", 1.0 if args.clip else 0.0)
    tb.set_camera_cols(cam)
    tb.set_option("eval_impl", float(args.eval_impl))
    tile_w, tile_h = 64, 32
    tb.set_tile_shard(rank, world, tile_w, tile_h)

    frame = torch.zeros(H, W, 4, device=dev)
    depth = torch.zeros(H, W, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(512 * 1024 * 1024 // 4, device=dev)  # 512 MB > 126 MB L2
    max_tiles = tb.tile_count(W, H, 0, world)
    packed = torch.zeros(max_tiles * tile_w * tile_h * 5, device=dev)
    gathered = torch.zeros(world * packed.numel(), device=dev) if world > 1 else None

    def step():
        tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
        launches = tb.frame_stats().n_kernel_launches
        if world > 1:
            tb.pack_tiles(W, H, rank, world, frame.data_ptr(), depth.data_ptr(), packed.data_ptr(), stream)
            dist.all_gather_into_tensor(gathered, packed)
            launches += 1
            if rank == 0:
                for r in range(1, world):
                    tb.unpack_tiles(W, H, r, world, gathered[r * packed.numel():].data_ptr(), frame.data_ptr(), depth.data_ptr(), stream)
                    launches += 1
        return launches

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    total_ms, launches = 0.0, 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    samples = 0
    for _ in range(args.steps):
        flush.fill_(1.0)  # L2 flush between timed iterations (untimed)
        barrier()
        e0.record()
        launches += step()
        e1.record()
        torch.cuda.synchronize()
        total_ms += e0.elapsed_time(e1)
        samples = tb.frame_stats().n_samples
    last_stats = tb.frame_stats()
    barrier()
    ms = total_ms / args.steps
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        s_t = torch.tensor([samples], device=dev, dtype=torch.int64)
        dist.all_reduce(s_t)
        samples_all = int(s_t.item())
    else:
        samples_all = int(samples)
    clocks = sampler.stop() if rank == 0 else None

    # ---- per-kernel durations (CUDA events on the launching stream, hn option "profile"), N=1 roofline of the dominant kernel
    tb.set_option("profile", 1.0)
    prof = {"ms_eval": 0.0, "ms_march": 0.0, "ms_composite": 0.0, "ms_mesh": 0.0, "ms_raygen": 0.0, "ms_envlight": 0.0, "n_passes": 0}
    n_prof = 3
    for _ in range(n_prof):
        flush.fill_(1.0)
        torch.cuda.synchronize()
        tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
        st = tb.frame_stats()
        for k in prof:
            prof[k] += getattr(st, k) / n_prof
    tb.set_option("profile", 0.0)
    hbm_peak, tf_peak, peak_kind = peaks()
    eval_s = prof["ms_eval"] * 1e-3
    ach_gbs = samples * ALGO_BYTES_PER_SAMPLE / eval_s / 1e9 if eval_s > 0 else 0.0
    roofline = {"kernel": "eval_kernel (fused hash-grid gather + density/colour MLPs)", "bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak,
                "unit": "GB/s", "frac": ach_gbs / hbm_peak, "traffic": NCU_EVAL_DRAM_BYTES,
                "traffic_source": "profiles/r01_ncu_eval_tc_v5.txt: dram read 142.8 MB + write 43.9 MB for the largest pass launch (727 us); far below the algorithmic gather bytes because the 23 MB of hash tables stay resident in the 126 MB L2 (88.9 % L2 hit rate) - compulsory DRAM bytes are the 32 B/sample of sample + output records",
                "peak_source": f"MEASURED_PEAKS.json ({peak_kind})",
                "mlp_tflops": samples * ALGO_FLOP_PER_SAMPLE / eval_s / 1e12 if eval_s > 0 else 0.0, "mlp_frac_of_tensor_peak": (samples * ALGO_FLOP_PER_SAMPLE / eval_s / 1e12) / tf_peak if eval_s > 0 else 0.0,
                "kernel_ms_per_frame": {k: round(v, 4) for k, v in prof.items()}, "samples_per_frame": int(samples)}

    # ---- e2e through Testbed.render(): camera from host, RGBA frame back to pinned host memory
    e2e_ms = None
    if world == 1:
        host_out = torch.empty(H, W, 4, pin_memory=True).numpy()
        tb.render(W, H, 1, True, out=host_out)
        t0 = time.perf_counter()
        n_e2e = max(3, min(args.steps, 10))
        for _ in range(n_e2e):
            tb.set_camera_cols(cam)
            tb.render(W, H, 1, True, out=host_out)
        e2e_ms = (time.perf_counter() - t0) / n_e2e * 1e3
    else:
        # multi-GPU e2e: the gathered frame is read back on rank 0
        host = torch.empty(H, W, 4, pin_memory=True) if rank == 0 else None
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(3, min(args.steps, 10))
        for _ in range(n_e2e):
            step()
            if rank == 0:
                host.copy_(frame, non_blocking=False)
            barrier()
        e2e_ms = (time.perf_counter() - t0) / n_e2e * 1e3
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.workload == "config3":
        import orc
        cores = os.cpu_count() or 1
        w8, h8 = W // 8, H // 8
        cfg = orc.net_config_from_info(info)
        bits, _ = orc.bitfield_from_grid(grid, 1)
        scene = orc.Scene([mesh])
        bb = scene.aabb()
        aabb = np.concatenate([bb[:3] - np.float32(4.0), bb[3:] + np.float32(4.0)])
        p = orc.render_params(w8, h8, cam, mesh_aabb=aabb, n_threads=cores, snap=True, rel_focal=tb.get_option("relative_focal_length", 2))
        params = hn.synthetic.random_params(info, 1337, 1e-4)
        t0 = time.perf_counter()
        orc.render_frame(p, cfg, params, bits, scene)
        dt = time.perf_counter() - t0
        cpu_baseline = {"value": w8 * h8 / dt / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
                        "sample": f"one {w8}x{h8} frame of the same view (1/64 of the 1080p rays), {dt:.1f} s"}

    if rank == 0:
        mrays = W * H / (ms * 1e-3) / 1e6
        workload = {"config2": "configs[1]: 1920x1080 NeRF-only (base.json, 128^3 sphere occupancy, random-init)",
                    "config3": "configs[2]: 1920x1080 hybrid NeRF (base.json, 128^3 sphere occupancy, random-init) + 1 048 576-triangle torus, sun shadow rays",
                    "config4": "configs[3]: 3840x2160 hybrid NeRF + 1 048 576-triangle torus, sun shadow rays, 1024x512 envmap background + "
                               "16 env light rays per mesh hit (BVH shadow ray + NeRF transmittance)"}[args.workload]
        line = {
            "metric": "Mrays/s, 1080p hybrid NeRF+mesh frame (primary rays)" if args.workload == "config3" else f"Mrays/s, {W}x{H} frame (primary rays), {args.workload}", "value": mrays, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "fp16 params/activations, fp32 accumulate", "data": "synthetic",
            "config": {"workload": workload,
                       "depth_merge": "clip" if args.clip else "reference per-pixel test", "l2": "512 MB flush between timed iterations",
                       "tiles": f"{tile_w}x{tile_h} interleaved over {world} rank(s)", "samples_per_frame_all_ranks": samples_all},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": W * H / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": 48, "d2h_bytes_per_step": W * H * 16},
            "roofline": roofline, "cpu_baseline": cpu_baseline,
        }
        if args.workload == "config4":
            line["env_light"] = {"hits_rank0": int(last_stats.n_env_hits), "light_rays_rank0": int(last_stats.n_light_rays),
                                 "light_rays_in_volume_rank0": int(last_stats.n_light_rays_volume), "light_samples_rank0": int(last_stats.n_light_samples)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
