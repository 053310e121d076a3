#!/usr/bin/env python
"""Reference-GPU baseline probe (measurement tooling, runs on the GPU box): times the reference's own CUDA build
(oracle/_ref/libngp_ref_gpu.so) and the product on the SAME scene files for BASELINE.json configs[1] (1080p NeRF-only)
and configs[2] (1080p hybrid, 1 Mi-triangle torus, sun shadow rays), and tcnn's network sequence on a sample stream.
Prints JSON lines; bench.py's reference arm uses the same calls."""
import argparse
import importlib
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refgpu  # noqa: E402

hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
syn = hn.synthetic


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nu", type=int, default=1024)
    ap.add_argument("--nv", type=int, default=512)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--w", type=int, default=1920)
    ap.add_argument("--h", type=int, default=1080)
    ap.add_argument("--skip-ours", action="store_true")
    args = ap.parse_args()
    import torch
    W, H = args.w, args.h
    tmp = tempfile.mkdtemp(prefix="refbench_")
    tb = hn.Testbed()
    tb.reload_network_from_file(os.path.join(ROOT, "configs", "nerf", "base.json"))
    info = tb.network_info()
    params = syn.random_params(info, 1337, 1e-4)
    grid = syn.sphere_density_grid(0.35)
    tb.set_params(params)
    tb.set_density_grid(grid, 1)
    tb.snap_to_pixel_centers = True
    cam = syn.default_camera()
    tb.set_camera_cols(cam)
    snap = os.path.join(tmp, "nerf.ingp")
    tb.save_snapshot(snap)
    rel_focal = tb.get_option("relative_focal_length", 2)

    def ours_ms(t, n=args.steps):
        frame = torch.zeros(H, W, 4, device="cuda"); depth = torch.zeros(H, W, device="cuda")
        s = torch.cuda.current_stream().cuda_stream
        flush = torch.empty(512 << 18, device="cuda")
        for _ in range(3):
            t.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, s)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot = 0.0
        for _ in range(n):
            flush.fill_(1.0); torch.cuda.synchronize()
            e0.record(); t.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, s); e1.record(); torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / n, int(t.frame_stats().n_samples)

    # ---- configs[1]: NeRF only
    rt = refgpu.RefTestbed()
    rt.load_snapshot(snap)
    rt.set("snap_to_pixel_centers", 1.0); rt.set_camera(cam); rt.set("relative_focal_length", rel_focal); rt.set("screen_center", [0.5, 0.5])
    ms = rt.time_render(W, H, 3, args.steps)
    line = {"what": "reference CUDA build, configs[1] NeRF-only", "res": [W, H], "ms_per_frame": float(ms.mean()), "ms_min": float(ms.min()), "Mrays_s": W * H / ms.mean() / 1e3}
    if not args.skip_ours:
        o_ms, o_s = ours_ms(tb)
        line.update({"ours_ms_per_frame": o_ms, "ours_samples": o_s, "speedup": float(ms.mean()) / o_ms})
    print(json.dumps(line, default=float), flush=True)
    # ---- tcnn network sequence on a sample stream of the same view: all steps of every ray through the occupancy sphere, step-major
    # (the reference's NerfCoordinate layout: sample j of ray i of a pass sits at i + j * n_alive)
    o, d, t0, alive = tb.op_init_rays(cam, cam, W, H, 0)
    o_t = torch.from_numpy(o).cuda(); d_t = torch.from_numpy(d).cuda()
    c = torch.tensor([0.5, 0.5, 0.5], device="cuda")
    oc = o_t - c
    b = (oc * d_t).sum(-1)
    disc = b * b - ((oc * oc).sum(-1) - 0.35 ** 2)
    hit = disc > 0
    sq = torch.sqrt(torch.clamp(disc, min=0))
    tn, tf = -b - sq, -b + sq
    dt = 1.7320508 / 1024
    nst = torch.where(hit, torch.clamp(((tf - tn) / dt).floor().long(), min=0), torch.zeros_like(b, dtype=torch.long))
    order = torch.argsort(nst, descending=True)
    nst_s = nst[order]
    kmax = int(nst_s[0].item())
    chunks = []
    for k in range(kmax):
        n_alive = int((nst_s > k).sum().item())
        if n_alive == 0:
            break
        idx = order[:n_alive]
        pos = o_t[idx] + (tn[idx] + (k + 0.5) * dt)[:, None] * d_t[idx]
        co = torch.cat([pos.clamp(0, 1), torch.full((n_alive, 1), dt, device="cuda"), (d_t[idx] + 1) * 0.5], 1)
        chunks.append(co)
    coords = torch.cat(chunks, 0).contiguous().float()
    n_total = coords.shape[0] // 256 * 256
    torch.cuda.synchronize()
    ms_net = rt.time_inference(coords.data_ptr(), n_total, 2 << 20, 3)
    print(json.dumps({"what": "tcnn NerfNetwork::inference_mixed_precision (grid + density MLP + SH + rgb MLP + extract_density), calls of 2 Mi samples",
                      "samples": n_total, "ms": ms_net, "Gsamples_s": n_total / ms_net / 1e6}, default=float), flush=True)
    del coords, chunks
    rt.close()

    # ---- configs[2]: hybrid, same scene files on both sides
    mesh = syn.torus_mesh(args.nu, args.nv)
    t0_ = time.perf_counter()
    obj = os.path.join(tmp, "torus.obj")
    refgpu.write_obj(obj, mesh)
    scene = os.path.join(tmp, "geometry_bench.json")
    refgpu.write_scene(scene, [("Mesh", obj, (0, 0, 0)), ("Nerf", snap, (0, 0, 0))])
    t1_ = time.perf_counter()
    rt = refgpu.RefTestbed()
    rt.load_training_data(scene)
    t2_ = time.perf_counter()
    rt.set("snap_to_pixel_centers", 1.0); rt.set("render_mode", 1.0); rt.set_camera(cam); rt.set("relative_focal_length", rel_focal); rt.set("screen_center", [0.5, 0.5])
    for label, crop in (("Geometry-mode default (render box = mesh box + 4)", None), ("render box cropped to the unit cube", [0, 0, 0, 1, 1, 1])):
        if crop is not None:
            rt.set("render_aabb", crop)
        ms = rt.time_render(W, H, 3, args.steps)
        line = {"what": "reference CUDA build, configs[2] hybrid", "variant": label, "triangles": int(mesh.shape[0]), "ms_per_frame": float(ms.mean()), "ms_min": float(ms.min()),
                "Mrays_s": W * H / ms.mean() / 1e3, "obj_write_s": t1_ - t0_, "ref_scene_load_s": t2_ - t1_}
        if not args.skip_ours:
            if crop is None:
                tb.load_file(scene)
                tb.set_camera_cols(cam); tb.snap_to_pixel_centers = True
                tb.set_option("compat_nearest_aabb_only", 1.0)
            else:
                tb.set_option("render_aabb", crop)
            o_ms, o_s = ours_ms(tb)
            line.update({"ours_ms_per_frame": o_ms, "ours_samples": o_s, "speedup": float(ms.mean()) / o_ms})
        print(json.dumps(line, default=float), flush=True)
    # parity of the two full-size frames (display space)
    disp = rt.render(W, H, 1, True)
    if not args.skip_ours:
        img = tb.render(W, H, 1, True)
        dd = np.abs(img - disp).max(-1)
        mse = float(np.mean((img.astype(np.float64) - disp) ** 2))
        print(json.dumps({"what": "1080p hybrid frame, ours vs reference CUDA build (unit-cube crop)", "max_abs": float(dd.max()), "pixels_gt_1e-2": int((dd > 1e-2).sum()),
                          "psnr_db": 10 * np.log10(1.0 / mse) if mse > 0 else 99.0}, default=float), flush=True)


if __name__ == "__main__":
    main()
