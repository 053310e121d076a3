#!/usr/bin/env python
"""Generates tests/golden/ref_gpu_*.npz: outputs of the REFERENCE'S OWN CUDA BUILD (oracle/_ref/libngp_ref_gpu.so: the
unmodified ngp::Testbed + tiny-cuda-nn sources of /root/reference compiled for sm_100a by oracle/ref_gpu.mk) on seeded
inputs, run on a B200. These pin the device-only half of the oracle (hash grid, SH, fused MLPs, compositing, mesh
tracing, whole frames), which the host-callable golden vectors (make_golden.py) cannot reach.

Needs a GPU: run on the GPU box, e.g.
  gpurun -- 'python tests/golden/make_ref_gpu_golden.py --out gpurun_out/refgpu'
then copy gpurun_out/refgpu/ref_gpu_*.npz to tests/golden/. Scene state goes INTO the reference through its own file
formats: the .ingp snapshot written by the product's hn_save_snapshot (read by Testbed::load_snapshot), ASCII .obj
meshes and the {"geometry": [...]} scene JSON (read by Testbed::load_scene) - so the run also witnesses format
compatibility. Every stage prints |ours - ref| and |oracle - ref| so drift is visible at generation time.
"""
import argparse
import importlib
import os
import sys
import tempfile
import traceback

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import orc  # noqa: E402
import refgpu  # noqa: E402

hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
syn = hn.synthetic
f32 = np.float32
TMP = tempfile.mkdtemp(prefix="refgpu_")


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


def ulp16(a, b):
    """distance in fp16 representable steps (sign-magnitude -> monotone integer)."""
    def key(x):
        u = x.view(np.uint16).astype(np.int32)
        return np.where(u & 0x8000, -(u & 0x7FFF), u)
    return np.abs(key(np.ascontiguousarray(a, np.float16)) - key(np.ascontiguousarray(b, np.float16)))


def make_product(cfg_rel, blob_range, aabb_scale=1, cascades=1):
    tb = hn.Testbed()
    if aabb_scale != 1:
        tb.set_aabb_scale(aabb_scale)
    tb.reload_network_from_file(os.path.join(ROOT, "configs", cfg_rel))
    info = tb.network_info()
    params = syn.random_params(info, 1337, blob_range)
    grid = syn.sphere_density_grid(0.35, cascades)
    tb.set_params(params)
    tb.set_density_grid(grid, cascades)
    tb.snap_to_pixel_centers = True
    tb.set_camera_cols(syn.default_camera())
    return tb, info, params, grid


def seeded_positions(rng, n):
    p = rng.random((n, 3), dtype=f32)
    corner = np.array([[0, 0, 0], [1, 1, 1], [0.999999, 0.5, 0.25], [1e-7, 1e-7, 1e-7], [0.5, 0.5, 0.5], [1, 0, 1], [0.25, 0.75, 1.0],
                       [0.0625, 0.03125, 0.015625]], f32)
    p[:corner.shape[0]] = corner
    # a run of consecutive ray steps (what a warp sees)
    o = np.array([0.31, 0.42, 0.93], f32); d = np.array([0.12, 0.05, -0.99], f32); d /= np.linalg.norm(d)
    k = min(256, n - 8)
    p[8:8 + k] = o + (np.arange(k, dtype=f32) * f32(1.7320508 / 1024))[:, None] * d
    return np.clip(p, 0, 1).astype(f32)


def stage_network(out_dir, tag, cfg_rel, blob_range, frame_res, aabb_scale=1, cascades=1):
    """encode / SH / network / NeRF-only frame for one (config, blob)."""
    tb, info, params, grid = make_product(cfg_rel, blob_range, aabb_scale, cascades)
    snap = os.path.join(TMP, f"{tag}.ingp")
    tb.save_snapshot(snap)
    rt = refgpu.RefTestbed()
    rt.load_snapshot(snap)
    rt.set("snap_to_pixel_centers", 1.0)
    rt.set_camera(syn.default_camera())
    rt.set("relative_focal_length", tb.get_option("relative_focal_length", 2))
    rt.set("screen_center", [0.5, 0.5])
    print(f"[{tag}] reference mode={rt.mode()} n_params={rt.get('n_params'):.0f} ours={info.n_params} widths={rt.widths()} "
          f"max_cascade={rt.get('max_cascade')} cone={rt.get('cone_angle_constant')} aabb={rt.get('aabb', 6)} render_aabb={rt.get('render_aabb', 6)}")
    res = {"level_scales": np.array([info.level_scales[i] for i in range(info.n_levels)], f32)}
    # parameters as the reference's trainer holds them, occupancy bits as its kernels built them from the snapshot's grid
    rp = rt.params(int(info.n_params))
    print(f"[{tag}] params bit-identical after the snapshot round trip: {np.array_equal(rp.view(np.uint16), params.view(np.uint16))}")
    rb = rt.bitfield()
    ob = tb.get_occupancy_bitfield()
    orb, _ = orc.bitfield_from_grid(grid, cascades)
    print(f"[{tag}] bitfield ref==ours {np.array_equal(rb, ob)}  ref==oracle {np.array_equal(rb, orb)}  set bits {int(np.unpackbits(rb).sum())}")
    res["bitfield_packed"] = np.packbits(np.unpackbits(rb))  # same bytes; stored whole (npz compresses the zeros)
    cfg = orc.net_config_from_info(info)
    rng = np.random.default_rng(20261105)
    n = 4096
    pos = seeded_positions(rng, n)
    # a-10 hash grid
    r_enc = rt.encode_pos(pos)
    o_enc = orc.encode(cfg, params, pos) if info.n_levels * info.n_features_per_level == 32 else None
    g_enc = tb.op_encode(pos)
    res["pos"], res["ref_encode"] = pos, r_enc
    print(f"[{tag}] encode: ours==ref bit-exact {np.array_equal(g_enc.view(np.uint16), r_enc.view(np.uint16))}, max ulp {int(ulp16(g_enc, r_enc).max())}; "
          + (f"oracle==ref {np.array_equal(o_enc.view(np.uint16), r_enc.view(np.uint16))}, max ulp {int(ulp16(o_enc, r_enc).max())}" if o_enc is not None else ""))
    # a-11 SH
    dirs = rng.normal(size=(n, 3)).astype(f32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(f32)
    dirs[:6] = [[1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [0, 0, -1]]
    d01 = ((dirs + f32(1.0)) * f32(0.5)).astype(f32)  # warp_direction
    r_sh = rt.encode_dir(d01)
    res["dir01"], res["ref_sh"] = d01, r_sh
    # a-12/13 network: samples in chunks of 32 per ray direction (the product's sample-stream shape)
    dt = np.full((n, 1), f32(1.7320508 / 1024), f32)
    dir_per_sample = np.repeat(d01[: n // 32], 32, axis=0)
    coords = np.concatenate([pos, dt, dir_per_sample], 1).astype(f32)
    r_net = rt.network(coords)
    o_net = orc.network(cfg, params, coords)
    g_net = tb.op_network(np.concatenate([pos, dt], 1), dirs[: n // 32])
    res["coords"], res["ref_network"] = coords, r_net
    for name, x in (("ours", g_net), ("oracle", o_net)):
        dd = np.abs(x.astype(f32) - r_net.astype(f32))
        rel = dd / (np.abs(r_net.astype(f32)) + 1e-3)
        print(f"[{tag}] network {name} vs ref: max abs {dd.max():.3e} (rgb {dd[:, :3].max():.3e}, sigma {dd[:, 3].max():.3e}), max rel {rel.max():.3e}")
    sh_o = None
    try:
        sh_o = orc.sh4(d01) if hasattr(orc, "sh4") else None
    except Exception:
        pass
    if sh_o is not None:
        print(f"[{tag}] SH oracle vs ref max ulp {int(ulp16(sh_o, r_sh).max())}")
    # whole NeRF-only frame (config 1 shape): raw frame buffer + depth of the reference
    if frame_res:
        w = h = frame_res
        disp, fr, dp = rt.render(w, h, 1, True, raw=True)
        img, depth = tb.render(w, h, 1, True, return_depth=True)
        oimg, odepth, osteps = orc.render_reference_frame(tb, params, grid, cascades, None, syn.default_camera(), w, h)
        res["frame_raw"], res["frame_depth"], res["frame_display"] = fr.astype(np.float16), dp, disp.astype(np.float16)
        res["frame_res"] = np.array([w, h])
        cov = float((fr[..., 3] > 0).mean())
        print(f"[{tag}] frame {w}x{h}: coverage {cov:.3f}; ours(display) vs ref(display) max {np.abs(img - disp).max():.3e} psnr {psnr(img, disp):.1f} dB; "
              f"oracle vs ref max {np.abs(oimg - disp).max():.3e} psnr {psnr(oimg, disp):.1f} dB; ours vs oracle max {np.abs(img - oimg).max():.3e}")
        m = (dp < 1000) & (depth < 1000)
        if m.any():
            print(f"[{tag}] depth: both-hit {int(m.sum())}, max rel diff {np.max(np.abs(dp[m] - depth[m]) / np.maximum(dp[m], 1e-3)):.3e}; "
                  f"hit-mask mismatch {int(((dp < 1000) != (depth < 1000)).sum())}")
    np.savez_compressed(os.path.join(out_dir, f"ref_gpu_{tag}.npz"), **res)
    rt.close()


def stage_hybrid(out_dir, tag, meshes, centers, res_wh, cam):
    """Geometry mode: meshes through .obj + scene JSON, NeRF through the .ingp snapshot entry."""
    tb, info, params, grid = make_product("nerf/base.json", 1.0)
    snap = os.path.join(TMP, f"{tag}.ingp")
    tb.save_snapshot(snap)
    entries = []
    for i, (m, c) in enumerate(zip(meshes, centers)):
        p = os.path.join(TMP, f"{tag}_{i}.obj")
        refgpu.write_obj(p, m)
        entries.append(("Mesh", p, c))
    entries.append(("Nerf", snap, (0, 0, 0)))
    scene = os.path.join(TMP, f"geometry_{tag}.json")
    refgpu.write_scene(scene, entries)
    rt = refgpu.RefTestbed()
    rt.load_training_data(scene)
    rt.set("snap_to_pixel_centers", 1.0)
    rt.set("render_mode", 1.0)
    rt.set_camera(cam)
    rt.set("relative_focal_length", tb.get_option("relative_focal_length", 2))
    rt.set("screen_center", [0.5, 0.5])
    rt.set("sun_dir", tb.get_option("sun_dir", 3))
    print(f"[{tag}] reference mode={rt.mode()} meshes={rt.mesh_count()} aabb={rt.get('aabb', 6)} render_aabb={rt.get('render_aabb', 6)} "
          f"nerf_bb={rt.get('nerf_bounding_box', 6)} sun={rt.get('sun_dir', 3)}")
    # product side through the SAME files
    tb.load_file(scene)
    tb.set_camera_cols(cam)
    tb.snap_to_pixel_centers = True
    w, h = res_wh
    res = {"cam": np.asarray(cam, f32), "res": np.array([w, h]), "centers": np.asarray(centers, f32)}
    # triangles as the reference holds them after load_mesh's normalisation (+ BVH reordering)
    ref_tris = [rt.mesh_triangles(i) for i in range(rt.mesh_count())]
    for i, t in enumerate(ref_tris):
        res[f"mesh_in_{i}"] = np.asarray(meshes[i], f32)
        res[f"ref_tris_{i}"] = t  # as the reference holds them: load_mesh-normalised, reordered by its BVH build
        key = lambda a: a[np.lexsort(a.T[::-1])]
        # GeometryBvh::build_mesh nth_element()s the mesh vector itself, so reference mesh i is SOME input mesh
        same = [np.array_equal(key(t).view(np.uint32), key(orc.normalize_mesh(m, c)).view(np.uint32)) for m, c in zip(meshes, centers)]
        print(f"[{tag}] reference mesh {i}: load_mesh-normalised triangles bit-identical (as a set) to the oracle's normalisation of input mesh {same.index(True) if any(same) else None}")
    # a-17/18 mesh trace on seeded rays (camera rays of this view + random + axis-aligned)
    rng = np.random.default_rng(7)
    o_c, d_c, _, _ = tb.op_init_rays(cam, cam, 64, 48, 0)
    n_r = 2048
    o_r = rng.uniform(-0.5, 1.5, (n_r, 3)).astype(f32)
    tgt = rng.uniform(0.2, 0.8, (n_r, 3)).astype(f32)
    d_r = tgt - o_r
    d_r /= np.linalg.norm(d_r, axis=1, keepdims=True).astype(f32)
    o_a = rng.uniform(0.1, 0.9, (256, 3)).astype(f32); o_a[:, 2] = 2.0
    d_a = np.tile(np.array([[0, 0, -1]], f32), (256, 1))
    ro = np.concatenate([o_c, o_r, o_a], 0); rd = np.concatenate([d_c, d_r, d_a], 0)
    pos_after, dir_after = rt.trace_mesh(ro, rd)
    res["trace_o"], res["trace_d"], res["ref_trace_pos"], res["ref_trace_dir"] = ro, rd, pos_after, dir_after
    mesh_i, tri_i, t_hit = tb.op_trace(ro, rd)
    t_ref = np.einsum("ij,ij->i", pos_after - ro, rd)
    changed = np.any(dir_after != rd, axis=1)
    print(f"[{tag}] trace: ref hits {int(changed.sum())}/{ro.shape[0]}, ours hits {int((tri_i >= 0).sum())}; on common hits max |t_ref - t_ours| = "
          f"{np.max(np.abs(t_ref - t_hit)[changed & (tri_i >= 0)]) if (changed & (tri_i >= 0)).any() else 0:.3e}")
    disp, fr, dp = rt.render(w, h, 1, True, raw=True)
    res["frame_raw"], res["frame_depth"], res["frame_display"] = fr.astype(np.float16), dp, disp.astype(np.float16)
    for compat in (1.0, 0.0):
        tb.set_option("compat_nearest_aabb_only", compat)
        img, depth = tb.render(w, h, 1, True, return_depth=True)
        print(f"[{tag}] hybrid frame {w}x{h} compat={int(compat)}: ours vs ref display max {np.abs(img - disp).max():.3e} psnr {psnr(img, disp):.1f} dB; "
              f"pixels > 1e-2: {int((np.abs(img - disp).max(-1) > 1e-2).sum())}")
    np.savez_compressed(os.path.join(out_dir, f"ref_gpu_{tag}.npz"), **res)
    rt.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "refgpu"))
    ap.add_argument("--stage", default="all")
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    stages = {
        "nerf_init": lambda: stage_network(args.out, "nerf_init", "nerf/base.json", 1e-4, 256),
        "nerf_contrast": lambda: stage_network(args.out, "nerf_contrast", "nerf/base.json", 1.0, 256),
        "geom_contrast": lambda: stage_network(args.out, "geom_contrast", "geometry/base.json", 1.0, 128),
        "hybrid_torus": lambda: stage_hybrid(args.out, "hybrid_torus", [syn.torus_mesh(64, 32)], [(0, 0, 0)], (192, 128), syn.default_camera()),
        "hybrid_three": lambda: stage_hybrid(args.out, "hybrid_three", [syn.torus_mesh(48, 24, seed=s) for s in (1, 2, 3)],
                                             [(0, 0, 0), (1, 1, 1), (-1, 1, 1)], (192, 128), syn.look_at_camera((0.5, 1.6, 4.5), (0.5, 1.0, 0.8))),
    }
    rc = 0
    for name, fn in stages.items():
        if args.stage not in ("all", name):
            continue
        try:
            fn()
            print(f"[{name}] done", flush=True)
        except Exception:
            rc = 1
            print(f"[{name}] FAILED", flush=True)
            traceback.print_exc()
    sys.exit(rc)


if __name__ == "__main__":
    main()
