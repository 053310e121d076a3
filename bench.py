#!/usr/bin/env python
"""Headline benchmark: Mrays/s (and ms/frame) of one 1920x1080 hybrid NeRF + 1 Mi-triangle-mesh frame with sun shadow rays
(BASELINE.json configs[2]) on N B200s, frame tile-sharded across ranks and gathered over NCCL.

  python bench.py [--gpus N] [--steps K] [--warmup W]           our CUDA path (one process per GPU under torchrun for N > 1)
  python bench.py --impl reference ...                          the REFERENCE'S OWN CUDA BUILD (oracle/_ref/libngp_ref_gpu.so: unmodified
                                                                ngp::Testbed + tiny-cuda-nn compiled for sm_100a, oracle/ref_gpu.mk) on the same
                                                                scene files; falls back to the CPU oracle port where that library is absent

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


def ncu_eval_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of the largest eval_tc_kernel launch, from the committed ncu --set full capture
    of the SHIPPED kernel variant (profiles/ncu_eval_tc_current.json, written by profiles/extract_ncu.py); None if absent."""
    p = os.path.join(ROOT, "profiles", "ncu_eval_tc_current.json")
    if not os.path.exists(p):
        return None, None
    d = json.load(open(p))
    return float(d["dram_bytes"]), d


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0))), "measured"
    return 6650.0, 1400.0, "fallback"


class quiet_stdout:
    """The reference library logs to stdout (tlog, std::cout); this line-oriented tool must print ONE JSON line there. Inside the
    block file descriptor 1 points at stderr."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)
        return False


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


def _load_synthetic():
    """synthetic.py without importing the package (the reference arm must not need the product's CUDA library)"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("hn_synthetic", os.path.join(ROOT, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "synthetic.py"))
    synthetic = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synthetic)
    return synthetic


WORKLOADS = {
    "config2": "configs[1]: 1920x1080 NeRF-only (base.json, 128^3 sphere occupancy, random-init)",
    "config3": "configs[2]: 1920x1080 hybrid NeRF (base.json, 128^3 sphere occupancy, random-init) + 1 048 576-triangle torus (load_mesh-normalised), sun shadow rays, "
               "Geometry-mode scene as Testbed::load_scene sets it up (render box = mesh box + 4, nearest-box mesh rule)",
    "config4": "configs[3]: 3840x2160 hybrid NeRF + 1 048 576-triangle torus, sun shadow rays, 1024x512 envmap background + "
               "16 env light rays per mesh hit (BVH shadow ray + NeRF transmittance)",
}


def reference_arm(args):
    """--impl reference. Rank 0 only. With oracle/_ref/libngp_ref_gpu.so present (built in the authoring container by oracle/ref_gpu.mk
    from the UNMODIFIED sources under /root/reference, shipped to the box like any built .so) the timed path is the reference's own
    CUDA build through its own public API: Testbed::load_training_data(scene JSON) / load_snapshot, then Testbed::render_frame per
    step (CUDA events on the Testbed's stream). Its inputs are files written in pure python (gzip+msgpack .ingp, ASCII .obj, scene
    JSON) - nothing of the product is on that path. Without the library: the scalar CPU oracle port on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import refgpu
    if args.workload in ("config2", "config3") and refgpu.available():
        with quiet_stdout():
            line = reference_arm_cuda(args, refgpu)
        print(json.dumps(line))
        return
    if args.workload == "config5" and refgpu.available():
        with quiet_stdout():
            line = reference_arm_train(args, refgpu)
        print(json.dumps(line))
        return
    return reference_arm_cpu(args)


def reference_arm_cuda(args, refgpu):
    import tempfile
    import orc  # layout arithmetic only (n_params of the config), not on the timed path
    synthetic = _load_synthetic()
    tmp = tempfile.mkdtemp(prefix="bench_ref_")
    cfg = orc.net_config(8, 4, 19, 16, float(np.exp(np.float32(np.log(np.float32(128.0)) / np.float32(7.0)))), 1, 2)

    class Info:
        n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = 8, 4, 1, 2
        n_params = orc.lib.orc_n_params(orc.C.byref(cfg))
        grid_param_offset = orc.lib.orc_grid_param_offset(orc.C.byref(cfg))

    params = synthetic.random_params(Info, 1337, 1e-4)
    grid = synthetic.sphere_density_grid(0.35)
    cam = synthetic.default_camera()
    snap = os.path.join(tmp, "nerf.ingp")
    refgpu.write_snapshot_py(snap, os.path.join(ROOT, "configs", "nerf", "base.json"), params, grid, 1, cam)
    t0 = time.perf_counter()
    rt = refgpu.RefTestbed()
    if args.workload == "config3":
        mesh = synthetic.torus_mesh(MESH_NU, MESH_NV)
        obj = os.path.join(tmp, "torus.obj")
        refgpu.write_obj(obj, mesh)
        scene = os.path.join(tmp, "geometry_bench.json")
        refgpu.write_scene(scene, [("Mesh", obj, (0, 0, 0)), ("Nerf", snap, (0, 0, 0))])
        rt.load_training_data(scene)     # Geometry mode: load_mesh + load_snapshot + GeometryBvh, render box = mesh box + 4
    else:
        rt.load_snapshot(snap)           # Nerf mode
    load_s = time.perf_counter() - t0
    rt.set("snap_to_pixel_centers", 1.0)
    rt.set("render_mode", 1.0)
    rt.set_camera(cam)
    rt.set("relative_focal_length", [synthetic.DEFAULT_REL_FOCAL] * 2)
    rt.set("screen_center", [0.5, 0.5])
    sampler = ClockSampler(0)
    sampler.start()
    ms_all = rt.time_render(W, H, args.warmup, args.steps, 512 << 20)
    clocks = sampler.stop()
    ms = float(ms_all.mean())
    # e2e: what pyngp.Testbed.render() does - render_frame + the frame copied to the host
    rt.render(W, H, 1, True)
    n_e2e = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        rt.set_camera(cam)
        rt.render(W, H, 1, True)
    e2e_ms = (time.perf_counter() - t0) / n_e2e * 1e3
    variants = {}
    if args.workload == "config3":  # the same scene with the render box cropped to the NeRF's unit cube (a GUI crop-box setting)
        rt.set("render_aabb", [0, 0, 0, 1, 1, 1])
        variants["render_aabb_unit_cube_ms"] = float(rt.time_render(W, H, 2, max(3, args.steps // 4), 512 << 20).mean())
    mrays = W * H / (ms * 1e-3) / 1e6
    line = {
        "impl": "reference", "metric": "Mrays/s, 1080p hybrid NeRF+mesh frame (primary rays)" if args.workload == "config3" else f"Mrays/s, {W}x{H} frame (primary rays), {args.workload}",
        "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "fp16 params/activations, fp16 accumulate (tcnn wmma)", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "l2": "512 MB memset between timed frames",
                   "device": "1 x B200 (the reference's Geometry mode is single-GPU; every N reports the same run)", "scene_load_s": load_s},
        "clocks": clocks,
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": 0, "kind": "reference",
                         "sample": f"the reference's own CUDA build (oracle/_ref/libngp_ref_gpu.so, unmodified sources, sm_100a) on the GPU, whole {W}x{H} frames through "
                                   f"Testbed::render_frame, {args.steps} steps; it has no CPU implementation of this path"},
        "e2e": {"value": W * H / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": W * H * 16},
        "variants": variants,
    }
    rt.close()
    return line


def reference_arm_train(args, refgpu):
    """--impl reference --workload config5: Testbed::train(2^18) of the reference CUDA build on the same synthetic dataset, written as
    transforms.json + PNG frames (pure python) and loaded by the reference's own loader; same network config file."""
    import tempfile
    synthetic = _load_synthetic()
    tmp = tempfile.mkdtemp(prefix="bench_ref_train_")
    imgs, cams, focal = synthetic.sphere_dataset(64, 256)
    refgpu.write_nerf_dataset(tmp, imgs, cams, focal)
    rt = refgpu.RefTestbed()
    rt.load_training_data(tmp)
    rt.reload_network(os.path.join(ROOT, "configs", "nerf", "base.json"))
    sampler = ClockSampler(0)
    sampler.start()
    ms, samples = rt.time_train(1 << 18, max(args.warmup, 8), args.steps)
    clocks = sampler.stop()
    val = samples / (ms * 1e-3) / 1e6
    rt.close()
    return {"impl": "reference", "metric": "Msamples/s, NeRF training step (forward + backward + Adam), target batch 2^18 samples per GPU", "value": val,
            "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": max(args.warmup, 8), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp16 params/activations (tcnn), fp32 master weights", "data": "synthetic",
            "config": {"workload": "configs[4]: Testbed::train(2^18) of the reference CUDA build, configs/nerf/base.json, the same 64 analytic views 256x256 as PNG + transforms.json",
                       "samples_per_step": samples, "device": "1 x B200 (the reference has no multi-GPU training)"},
            "clocks": clocks, "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": 0, "kind": "reference",
                                               "sample": f"the reference's own CUDA build on the GPU, {args.steps} training steps (wall clock around Testbed::train)"},
            "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4}}


def reference_arm_cpu(args):
    """Fallback of --impl reference where the reference CUDA build is absent: the scalar host oracle (restatement of the reference
    pipeline, oracle/) on all host cores. Each step renders a bounded sample: the same camera/scene at reduced resolution."""
    import orc
    synthetic = _load_synthetic()
    cores = os.cpu_count() or 1
    w, h = W // 8, H // 8
    cfg = orc.net_config(8, 4, 19, 16, float(np.exp(np.float32(np.log(np.float32(128.0)) / np.float32(7.0)))), 1, 2)

    class Info:  # shape of hn_network_info for synthetic.random_params
        n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = 8, 4, 1, 2
        n_params = orc.lib.orc_n_params(orc.C.byref(cfg))
        grid_param_offset = orc.lib.orc_grid_param_offset(orc.C.byref(cfg))

    params = synthetic.random_params(Info, 1337, 1e-4)
    bits, _ = orc.bitfield_from_grid(synthetic.sphere_density_grid(0.35), 1)
    mesh = orc.normalize_mesh(synthetic.torus_mesh(MESH_NU, MESH_NV))
    scene = orc.Scene([mesh])
    bb = scene.aabb()
    aabb = np.concatenate([bb[:3] - np.float32(4.0), bb[3:] + np.float32(4.0)])
    # bound the whole run (warm-up + steps) to about 150 s of CPU work whatever K and W are: one 96x54 calibration frame gives
    # the host's ray rate, the per-step sample is the same view at the 16:9 resolution that fits the budget (at most 240x135)
    pc = orc.render_params(96, 54, synthetic.default_camera(), mesh_aabb=aabb, render_aabb=aabb, compat=True, n_threads=cores, snap=True)
    t0 = time.perf_counter()
    orc.render_frame(pc, cfg, params, bits, scene)
    rate = 96 * 54 / (time.perf_counter() - t0)
    rays = min(w * h, max(64 * 36, rate * 150.0 / max(1, args.warmup + args.steps)))
    h = max(36, int((rays * 9 / 16) ** 0.5))
    w = h * 16 // 9
    p = orc.render_params(w, h, synthetic.default_camera(), mesh_aabb=aabb, render_aabb=aabb, compat=True, n_threads=cores, snap=True)
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
        "config": {"workload": WORKLOADS["config3"],
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
        with quiet_stdout():  # NCCL announces its version on stdout when the first communicator comes up
            dist.init_process_group("nccl", device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize()
    hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
    tb = hn.Testbed(device=local)
    tb.reload_network_from_file(os.path.join(ROOT, "configs", "nerf", "base.json"))
    info = tb.network_info()
    tb.set_params(hn.synthetic.random_params(info, 1337, 1e-4))
    tb.set_density_grid(hn.synthetic.sphere_density_grid(0.35), 1)
    imgs, cams, focal = hn.synthetic.sphere_dataset(64, 256)
    tb.set_training_data(imgs, cams, focal)
    tb.set_tile_shard(rank, world)
    for kv in args.opt:
        k, v = kv.split("=")
        tb.set_option(k, [float(x) for x in v.split(",")])
    tb.reset_trainer(1337)
    ptr, n = tb.train_gradient_buffer()
    grads = torch.as_tensor(_DevArray(ptr, n), device=dev)

    if world > 1:  # the library's training calls go onto torch's current stream, the one NCCL's all-reduce is enqueued on:
        tb.set_stream(torch.cuda.current_stream().cuda_stream)  # accumulate -> all-reduce -> apply are stream-ordered, no host sync

    def step():
        loss = tb.train_accumulate(0)
        if world > 1:
            dist.all_reduce(grads)
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
    roofline, cpu_baseline = None, None
    if rank == 0:
        # roofline of the step's HBM-bound kernel: adam_kernel streams, per parameter, fp32 master (r+w), two fp32 moments (r+w), the fp32
        # gradient (r), the per-parameter step counter (r+w), the fp16 training copy (w) and the fp16 EMA copy + fp32 EMA accumulator (r+w / w)
        n_par = int(info.n_params)
        bytes_adam = n_par * (8 + 8 + 8 + 4 + 8 + 2 + 2 + 8)
        hbm_peak, tf_peak, peak_kind = peaks()
        if st.ms_optimizer > 0:
            ach = bytes_adam / (st.ms_optimizer * 1e-3) / 1e9
            roofline = {"kernel": "adam_kernel (Adam + EMA over all 11.7 M parameters; the step's bandwidth-bound kernel)", "bound": "hbm", "achieved": ach, "peak": hbm_peak,
                        "unit": "GB/s", "frac": ach / hbm_peak, "traffic": None, "peak_source": f"MEASURED_PEAKS.json ({peak_kind})",
                        "note": "the longest kernel of the step is train_march_kernel (one DDA walk per ray, latency-bound: ~6 800 rays x ~37 samples, no bandwidth or "
                                "tensor roofline applies); the three tcgen05 MLP kernels together take forward + backward ms below",
                        "mlp_tflops": 3 * st.n_samples_last * 18816 / ((st.ms_forward + st.ms_backward) * 1e-3) / 1e12 if st.ms_forward + st.ms_backward > 0 else None}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import orc
        cores = os.cpu_count() or 1
        cfg = orc.net_config_from_info(info)
        grid = hn.synthetic.sphere_density_grid(0.35)
        bits, _ = orc.bitfield_from_grid(grid, 1)
        rng = np.random.default_rng(5)
        nr = 96
        o = (0.5 + rng.normal(size=(nr, 3)) / np.linalg.norm(rng.normal(size=(nr, 3)), axis=1, keepdims=True) * 1.6).astype(np.float32)
        d = (0.5 + (rng.random((nr, 3)) - 0.5) * 0.5 - o).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        t0v = np.full(nr, 0.8, np.float32)
        p = orc.render_params(8, 8, hn.synthetic.default_camera(), n_threads=cores)
        tt = time.perf_counter()
        _, _, ws = orc.train_gradients(p, cfg, hn.synthetic.random_params(info, 1337, 1e-4), bits, o, d, t0v, rng.random((nr, 3), dtype=np.float32), rng.random((nr, 3), dtype=np.float32))
        dtc = time.perf_counter() - tt
        cpu_baseline = {"value": float(ws.sum()) / dtc / 1e6, "unit": "Msamples/s", "cores": cores, "kind": "port",
                        "sample": f"oracle forward + fp32 backward of {nr} rays = {int(ws.sum())} samples, {dtc:.1f} s (no optimizer step)"}
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
                "step_phases_ms": phases, "roofline": roofline, "cpu_baseline": cpu_baseline, "options": args.opt}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def reference_network_leg(tb, hn, cam, samples_frame, ms_eval, torch):
    """tcnn's NerfNetwork::inference_mixed_precision (kernel_grid x L, density kernel_mlp_fused, kernel_sh, rgb kernel_mlp_fused,
    extract_density: nerf_network.h:105-139) timed with CUDA events on a sample stream of the benched view - every step of every ray
    through the occupancy sphere, in the reference's step-major NerfCoordinate order, fed in calls of 2 Mi samples (the tracer's pass
    size, src/testbed_nerf.cu:2080-2081) - next to this project's eval_tc_kernel time for the frame's own stream."""
    import tempfile
    import refgpu
    if not refgpu.available():
        return {"unavailable": "oracle/_ref/libngp_ref_gpu.so not present on this box"}
    snap = os.path.join(tempfile.mkdtemp(prefix="bench_net_"), "nerf.ingp")
    tb.save_snapshot(snap)
    rt = refgpu.RefTestbed()
    rt.load_snapshot(snap)
    o, d, _, _ = tb.op_init_rays(cam, cam, W, H, 0)
    o_t, d_t = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    oc = o_t - 0.5
    b = (oc * d_t).sum(-1)
    disc = b * b - ((oc * oc).sum(-1) - 0.35 ** 2)
    sq = torch.sqrt(torch.clamp(disc, min=0))
    tn, tf = -b - sq, -b + sq
    dt = 1.7320508 / 1024
    nst = torch.where(disc > 0, torch.clamp(((tf - tn) / dt).floor().long(), min=0), torch.zeros_like(b, dtype=torch.long))
    order = torch.argsort(nst, descending=True)
    nst_s = nst[order]
    chunks = []
    for k in range(int(nst_s[0].item())):
        n_alive = int((nst_s > k).sum().item())
        idx = order[:n_alive]
        pos = o_t[idx] + (tn[idx] + (k + 0.5) * dt)[:, None] * d_t[idx]
        chunks.append(torch.cat([pos.clamp(0, 1), torch.full((n_alive, 1), dt, device="cuda"), (d_t[idx] + 1) * 0.5], 1))
    coords = torch.cat(chunks, 0).contiguous().float()
    del chunks
    n_total = coords.shape[0] // 256 * 256
    torch.cuda.synchronize()
    ms_net = rt.time_inference(coords.data_ptr(), n_total, 2 << 20, 3)
    rt.close()
    ours_gs = samples_frame / ms_eval / 1e6 if ms_eval > 0 else 0.0
    return {"what": "reference CUDA build (unmodified tiny-cuda-nn, sm_100a): NerfNetwork::inference_mixed_precision, 5 + L launches per call of 2 Mi samples",
            "samples": int(n_total), "ms": ms_net, "gsamples_s": n_total / ms_net / 1e6,
            "ms_per_frame_equivalent": ms_net * samples_frame / n_total,
            "ours_eval_tc_kernel": {"samples": int(samples_frame), "ms": ms_eval, "gsamples_s": ours_gs},
            "speedup_network": ours_gs / (n_total / ms_net / 1e6)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--eval-impl", type=int, default=1, help="network kernel: 1 = tcgen05/TMEM (default), 0 = warp MMA")
    ap.add_argument("--clip", action="store_true", help="stop marching at the mesh depth instead of the reference's per-pixel test")
    ap.add_argument("--crop-unit-cube", action="store_true", help="render box = the NeRF's unit cube instead of the Geometry-mode default (mesh box + 4)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"], help="multi-GPU frame gather: peer-memory tile scatter (default) or the round-1 NCCL all-gather")
    ap.add_argument("--opt", action="append", default=[], help="name=value library option (hn_set_option) for A/B runs, e.g. --opt eval_gather=2")
    ap.add_argument("--e2e-delivery", default="host", choices=["host", "gather"], help="N>1 e2e: tiles scattered straight to a shared host frame / gathered on rank 0 then copied")
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
        with quiet_stdout():  # NCCL announces its version on stdout when the first communicator comes up
            dist.init_process_group("nccl", device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize()

    hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
    mesh, grid, cam = build_scene_inputs(hn)
    tb = hn.Testbed(device=local)
    for kv in args.opt:  # options that act at scene-build time (bvh_sah) are set before the scene is built; all are applied again below
        k, v = kv.split("=")
        tb.set_option(k, [float(x) for x in v.split(",")])
    tb.reload_network_from_file(os.path.join(ROOT, "configs", "nerf", "base.json"))
    info = tb.network_info()
    tb.set_params(hn.synthetic.random_params(info, 1337, 1e-4))
    tb.set_density_grid(grid, 1)
    if args.workload != "config2":
        # the scene as Testbed::load_scene builds it from an .obj: load_mesh normalisation, render box = mesh root box + 4, nearest-box
        # mesh rule (identical arithmetic to loading the same files: tests/test_gpu_refgpu.py::test_live_hybrid_scene_files)
        tb.add_mesh(mesh, normalize=True)
        tb.build_scene()
        if not args.crop_unit_cube:
            tb.set_option("render_aabb", tb.get_option("mesh_aabb", 6))
        tb.set_option("compat_nearest_aabb_only", 1.0)
    if args.workload == "config4":
        tb.set_envmap(hn.synthetic.envmap(1024, 512, seed=3))
        tb.set_option("env_light_samples", 16.0)
    tb.snap_to_pixel_centers = True
    tb.set_option("depth_merge_clip", 1.0 if args.clip else 0.0)
    tb.set_camera_cols(cam)
    tb.set_option("eval_impl", float(args.eval_impl))
    for kv in args.opt:
        k, v = kv.split("=")
        tb.set_option(k, [float(x) for x in v.split(",")])
    tile_w, tile_h = 64, 32
    tb.set_tile_shard(rank, world, tile_w, tile_h)

    stream = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(512 * 1024 * 1024 // 4, device=dev)  # 512 MB > 126 MB L2
    # ---- multi-GPU exchange: "peer" (default) = every rank copies its tiles straight into rank 0's frame / depth over NVLink peer
    # memory (CUDA IPC mapping, hn_scatter_tiles: one launch) and one 4-byte all-reduce orders it; "nccl" = round-1 path
    # (pack -> all_gather of all 41.5 MB to every rank -> one unpack launch per peer on rank 0), kept for A/B
    exchange = args.exchange if world > 1 else "none"
    remote_frame = remote_depth = 0
    if exchange == "peer":
        handles = [None, None]
        if rank == 0:
            fptr, fh = tb.shared_alloc(W * H * 16)
            dptr, dh = tb.shared_alloc(W * H * 4)
            handles = [fh, dh]
            frame = torch.as_tensor(_DevArray(fptr, W * H * 4), device=dev).view(H, W, 4)
            depth = torch.as_tensor(_DevArray(dptr, W * H), device=dev).view(H, W)
        dist.broadcast_object_list(handles, src=0)
        if rank != 0:
            remote_frame, remote_depth = tb.shared_open(handles[0]), tb.shared_open(handles[1])
            frame = torch.zeros(H, W, 4, device=dev)
            depth = torch.zeros(H, W, device=dev)
        token = torch.zeros(1, device=dev)
    else:
        frame = torch.zeros(H, W, 4, device=dev)
        depth = torch.zeros(H, W, device=dev)
    if exchange == "nccl":
        max_tiles = tb.tile_count(W, H, 0, world)
        packed = torch.zeros(max_tiles * tile_w * tile_h * 5, device=dev)
        gathered = torch.zeros(world * packed.numel(), device=dev)

    def do_exchange():
        n = 0
        if exchange == "peer":
            if rank != 0:
                tb.scatter_tiles(W, H, rank, world, frame.data_ptr(), depth.data_ptr(), remote_frame, remote_depth, stream)
                n += 1
            dist.all_reduce(token)  # stream-ordered on every rank: when it completes on rank 0 every peer's tile stores have landed
        elif exchange == "nccl":
            tb.pack_tiles(W, H, rank, world, frame.data_ptr(), depth.data_ptr(), packed.data_ptr(), stream)
            dist.all_gather_into_tensor(gathered, packed)
            n += 1
            if rank == 0:
                for r in range(1, world):
                    tb.unpack_tiles(W, H, r, world, gathered[r * packed.numel():].data_ptr(), frame.data_ptr(), depth.data_ptr(), stream)
                    n += 1
        return n

    def step():
        tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
        return tb.frame_stats().n_kernel_launches + do_exchange()

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
    # ---- where a multi-GPU step goes: this rank's kernels (CUDA events per kernel), the exchange alone (entered together), the rest
    breakdown = None
    if world > 1:
        ex_ms = 0.0
        for _ in range(5):
            barrier()
            e0.record(); do_exchange(); e1.record(); torch.cuda.synchronize()
            ex_ms += e0.elapsed_time(e1) / 5
        kern = sum(v for k, v in prof.items() if k.startswith("ms_"))
        t = torch.tensor([kern, ex_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        kern, ex_ms = float(t[0].item()), float(t[1].item())
        # the gathered frame on rank 0 must equal the frame one GPU renders alone (tiles are independent): checked, untimed
        step(); barrier()
        gather_err = None
        if rank == 0:
            gathered_frame = frame.clone()
            tb.set_tile_shard(0, 1, tile_w, tile_h)
            full = torch.zeros(H, W, 4, device=dev); full_d = torch.zeros(H, W, device=dev)
            tb.render_frame(cam, cam, W, H, full.data_ptr(), full_d.data_ptr(), 0, stream)
            torch.cuda.synchronize()
            gather_err = float((gathered_frame - full).abs().max().item())
            tb.set_tile_shard(rank, world, tile_w, tile_h)
        barrier()
        breakdown = {"gathered_frame_max_abs_vs_unsharded_render": gather_err, "kernel_ms_per_frame": kern, "exchange_ms": ex_ms, "host_gap_ms": max(0.0, ms - kern - ex_ms), "exchange": exchange,
                     "note": "max over ranks; kernel = sum of the slowest rank's own kernels, exchange = tile scatter + ordering all-reduce timed alone after a barrier, "
                             "host_gap = step time minus both (launch gaps, the per-pass counter read-backs, rank skew)"}
    hbm_peak, tf_peak, peak_kind = peaks()
    eval_s = prof["ms_eval"] * 1e-3
    ach_gbs = samples * ALGO_BYTES_PER_SAMPLE / eval_s / 1e9 if eval_s > 0 else 0.0
    traffic, ncu = ncu_eval_traffic()
    roofline = {"kernel": "eval_tc_kernel (fused hash-grid gather + density/colour MLPs, tcgen05/TMEM)", "bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak,
                "unit": "GB/s", "frac": ach_gbs / hbm_peak, "traffic": traffic,
                "note": "achieved = ALGORITHMIC gather bytes (512 B/sample, SURVEY 8d) / kernel time: an effective gather rate against the measured HBM copy "
                        "bandwidth, as SURVEY 8(d) defines it. The 23 MB table is L2-resident, so the kernel's real DRAM traffic (`traffic`, per largest launch, from the "
                        "ncu capture of the shipped variant) is a few % of that and DRAM is NOT the binding unit: ncu names issue slots / the L1TEX gather path (see `ncu`)",
                "ncu": ncu, "peak_source": f"MEASURED_PEAKS.json ({peak_kind})",
                "mlp_tflops": samples * ALGO_FLOP_PER_SAMPLE / eval_s / 1e12 if eval_s > 0 else 0.0, "mlp_frac_of_tensor_peak": (samples * ALGO_FLOP_PER_SAMPLE / eval_s / 1e12) / tf_peak if eval_s > 0 else 0.0,
                "kernel_ms_per_frame": {k: round(v, 4) for k, v in prof.items()}, "samples_per_frame": int(samples)}

    # ---- the reference's own tcnn network sequence on the same box (N = 1; oracle/_ref/libngp_ref_gpu.so when it is present)
    reference_cuda = None
    if rank == 0 and world == 1 and args.workload in ("config2", "config3"):
        try:
            with quiet_stdout():
                reference_cuda = reference_network_leg(tb, hn, cam, samples, prof["ms_eval"], torch)
        except Exception as e:  # a baseline leg never takes the bench line down
            reference_cuda = {"unavailable": str(e)[:200]}

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
        # multi-GPU e2e: the frame is delivered to HOST memory. "host" (default): one shared-memory frame every rank has mapped and
        # page-locked (hn_host_register); each rank scatters the tiles it owns straight into it over its own PCIe link (hn_scatter_tiles,
        # one launch), then stream sync + barrier - no gather on rank 0, no 33 MB read-back through one link. "gather": the round-1 path
        # (peer gather into rank 0's device frame, then rank 0 copies it to pinned host memory), kept for A/B.
        n_e2e = max(3, min(args.steps, 10))
        delivery = args.e2e_delivery
        shared, host_dev = None, 0
        if delivery == "host":
            # set-up may fail on a box whose /dev/shm is too small or that refuses to page-lock the mapping: every rank then agrees to
            # fall back to the gather path (and the line says so)
            from importlib import import_module
            tiles_mod = import_module(hn.__name__ + ".tiles")
            ok = 1.0
            names = [None]
            if rank == 0:
                try:
                    names = [f"hn_e2e_{os.getpid()}"]
                    shared = tiles_mod.SharedHostFrame(names[0], H, W, 4, create=True)
                except Exception as e:
                    print(f"[bench] shared host frame unavailable: {e}", file=sys.stderr)
                    names, ok = [None], 0.0
            dist.broadcast_object_list(names, src=0)
            try:
                if names[0] is None:
                    ok = 0.0
                else:
                    if rank != 0:
                        shared = tiles_mod.SharedHostFrame(names[0], H, W, 4, create=False)
                    host_dev = tb.host_register(shared.host_ptr, shared.n_bytes)
            except Exception as e:
                print(f"[bench] rank {rank}: host delivery set-up failed: {e}", file=sys.stderr)
                ok = 0.0
            flag = torch.tensor([ok], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if float(flag.item()) < 1.0:
                if host_dev:
                    tb.host_unregister(shared.host_ptr)
                if shared is not None:
                    shared.close()
                shared, host_dev, delivery = None, 0, "gather"
        if delivery == "host":
            def e2e_step():
                tb.set_camera_cols(cam)
                tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
                tb.scatter_tiles(W, H, rank, world, frame.data_ptr(), 0, host_dev, 0, stream)
                torch.cuda.synchronize()
                dist.barrier()
        else:
            host = torch.empty(H, W, 4, pin_memory=True) if rank == 0 else None

            def e2e_step():
                step()
                if rank == 0:
                    host.copy_(frame, non_blocking=False)
                barrier()
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        e2e_ms = (time.perf_counter() - t0) / n_e2e * 1e3
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
        if breakdown is not None:
            breakdown["e2e_delivery"] = "peer gather on rank 0, then one device -> host copy there"
        if delivery == "host":
            # the host frame must be the frame the peer gather assembles on rank 0 (untimed check)
            step(); barrier()
            if rank == 0:
                e2e_check = float(np.abs(shared.array - frame.cpu().numpy()).max())
                if breakdown is not None:
                    breakdown["e2e_host_frame_max_abs_vs_gathered_frame"] = e2e_check
                    breakdown["e2e_delivery"] = "every rank scatters its tiles into one page-locked shared host frame (hn_host_register + hn_scatter_tiles)"
            barrier()
            tb.host_unregister(shared.host_ptr)
            barrier()
            shared.close()

    cpu_baseline, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.workload == "config3":
        # The CPU leg doubles as the parity check of the BENCHED configuration: the oracle renders a stratified 1/64 of the pixels of
        # this very 1920x1080 frame (every 8th row and column, same scene, same boxes), timed on all host cores, and the frame the
        # GPU just produced is compared at those pixels.
        import orc
        cores = os.cpu_count() or 1
        cfg = orc.net_config_from_info(info)
        bits, _ = orc.bitfield_from_grid(grid, 1)
        nmesh = orc.normalize_mesh(mesh)
        scene = orc.Scene([nmesh])
        bb = scene.aabb()
        aabb = np.concatenate([bb[:3] - np.float32(4.0), bb[3:] + np.float32(4.0)])
        mask = np.zeros((H, W), np.uint8)
        mask[::8, ::8] = 1
        n_px = int(mask.sum())
        rbox = tb.get_option("render_aabb", 6)
        p = orc.render_params(W, H, cam, mesh_aabb=aabb, render_aabb=rbox, compat=True, clip=args.clip, n_threads=cores, snap=True,
                              rel_focal=tb.get_option("relative_focal_length", 2), pixel_mask=mask)
        params = hn.synthetic.random_params(info, 1337, 1e-4)
        t0 = time.perf_counter()
        o_frame, o_depth, o_steps, _ = orc.render_frame(p, cfg, params, bits, scene)
        dt = time.perf_counter() - t0
        cpu_baseline = {"value": n_px / dt / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
                        "sample": f"{n_px} pixels of this {W}x{H} frame (every 8th row and column = 1/64 of the rays), {dt:.1f} s"}
        tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
        torch.cuda.synchronize()
        g_frame = frame.cpu().numpy()
        m = mask.astype(bool)
        diff = np.abs(g_frame[m] - o_frame[m])
        mse = float(np.mean((g_frame[m].astype(np.float64) - o_frame[m]) ** 2))
        o_, d_, t_, alive_ = tb.op_init_rays(cam, cam, W, H, 0)
        counts, _ = tb.op_march_all(o_[m.ravel()], d_[m.ravel()], t_[m.ravel()], alive_[m.ravel()])
        parity = {"against": "CPU oracle (pinned on the reference CUDA build, tests/test_ref_gpu_pin.py)", "pixels": n_px, "max_abs": float(diff.max()),
                  "pixels_gt_1e-2": int((diff.max(-1) > 1e-2).sum()), "psnr_db": 10 * np.log10(1.0 / mse) if mse > 0 else 99.0,
                  "per_ray_sample_counts_equal": bool(np.array_equal(counts, o_steps[m])),
                  "samples_oracle_x64": int(o_steps[m].sum()) * 64, "samples_gpu_frame": int(samples)}

    if rank == 0:
        mrays = W * H / (ms * 1e-3) / 1e6
        workload = WORKLOADS[args.workload]
        line = {
            "metric": "Mrays/s, 1080p hybrid NeRF+mesh frame (primary rays)" if args.workload == "config3" else f"Mrays/s, {W}x{H} frame (primary rays), {args.workload}", "value": mrays, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "fp16 params/activations, fp32 accumulate", "data": "synthetic",
            "config": {"workload": workload,
                       "depth_merge": "clip" if args.clip else "reference per-pixel test", "l2": "512 MB flush between timed iterations",
                       "tiles": f"{tile_w}x{tile_h} interleaved over {world} rank(s)", "samples_per_frame_all_ranks": samples_all, "options": args.opt},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": W * H / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": 48, "d2h_bytes_per_step": W * H * 16},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "parity": parity, "reference_cuda": reference_cuda, "multi_gpu": breakdown,
        }
        if args.workload == "config4":
            line["env_light"] = {"hits_rank0": int(last_stats.n_env_hits), "light_rays_rank0": int(last_stats.n_light_rays),
                                 "light_rays_in_volume_rank0": int(last_stats.n_light_rays_volume), "light_samples_rank0": int(last_stats.n_light_samples)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _exit_quietly():
    """The reference library's logger writes an ANSI reset to stdout from a static destructor at process exit, i.e. AFTER the JSON line.
    When that library was loaded the process therefore leaves through os._exit once its own output is flushed."""
    mod = sys.modules.get("refgpu")
    if mod is not None and getattr(mod, "_lib", None) is not None:
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
    _exit_quietly()
