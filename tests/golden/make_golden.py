#!/usr/bin/env python
"""Generates the committed golden vectors under tests/golden/.

  ref_host.npz   outputs of the REFERENCE's own NGP_HOST_DEVICE functions (compiled for the host from /root/reference by
                 `make -C oracle ref` -> oracle/_ref/libngp_ref_host.so) on seeded inputs. Needs /root/reference, so it
                 can only be produced in the authoring container; the tests read the .npz and never the reference.
  oracle_pin.npz outputs of the CPU restatement (oracle/liborc.so) for the device-only kernels the reference cannot run on a
                 host (hash grid, SH, MLP, compositing, shading, whole frames). The reference ships no vectors for them
                 (SURVEY.md §4), so these pin the oracle against accidental change only ("parity unpinned" by the reference).

Run: python tests/golden/make_golden.py   (after `python -c "import __graft_entry__ as g; g.build()"`)
"""
import ctypes as C
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import orc  # noqa: E402

spec = importlib.util.spec_from_file_location("hn_synthetic", os.path.join(ROOT, "surface-irradiance-estimation-from-neural-radiance-fields_b200", "synthetic.py"))
synthetic = importlib.util.module_from_spec(spec)
spec.loader.exec_module(synthetic)

f32 = np.float32
P = orc.P


def fa(*v):
    return (C.c_float * len(v))(*[float(x) for x in v])


def make_ref_host():
    ref = orc.ref
    if ref is None:
        raise SystemExit("oracle/_ref/libngp_ref_host.so missing: run `make -C oracle ref` where /root/reference is mounted")
    rng = np.random.default_rng(20261104)
    out = {}
    # a-2 Sobol / Owen scrambling
    idx = rng.integers(0, 2 ** 32, 256, dtype=np.uint64).astype(np.uint32)
    seed = rng.integers(0, 2 ** 32, 256, dtype=np.uint64).astype(np.uint32)
    idx[:8] = [0, 1, 2, 3, 4, 5, 6, 7]
    out["ld_idx"], out["ld_seed"] = idx, seed
    out["ld_val"] = np.array([ref.ref_ld_random_val(C.c_uint32(int(i)), C.c_uint32(int(s))) for i, s in zip(idx, seed)], f32)
    po = np.zeros((32, 2), f32)
    for spp in range(32):
        tmp = (C.c_float * 2)()
        ref.ref_ld_random_pixel_offset(C.c_uint32(spp), tmp)
        po[spp] = tmp[:]
    out["pixel_offset"] = po
    # a-5 slab test, a-17 triangle test
    n = 512
    bb = np.concatenate([rng.uniform(-1, 0.4, (n, 3)), rng.uniform(0.6, 2, (n, 3))], 1).astype(f32)
    o = rng.uniform(-2, 3, (n, 3)).astype(f32)
    d = rng.normal(size=(n, 3)).astype(f32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(f32)
    d[:4, 0] = 0.0  # axis-parallel rays: division by zero inside the slab test
    tt = np.zeros((n, 2), f32)
    for i in range(n):
        tmp = (C.c_float * 2)()
        ref.ref_bbox_ray_intersect(fa(*bb[i]), fa(*o[i]), fa(*d[i]), tmp)
        tt[i] = tmp[:]
    out["bb"], out["bb_o"], out["bb_d"], out["bb_t"] = bb, o, d, tt
    tri = rng.uniform(0, 1, (n, 9)).astype(f32)
    to = rng.uniform(-1, 2, (n, 3)).astype(f32)
    tgt = (tri[:, 0:3] + tri[:, 3:6] + tri[:, 6:9]) / f32(3) + rng.normal(scale=0.15, size=(n, 3)).astype(f32)
    td = (tgt - to).astype(f32)
    td /= np.linalg.norm(td, axis=1, keepdims=True).astype(f32)
    out["tri"], out["tri_o"], out["tri_d"] = tri, to, td
    out["tri_t"] = np.array([ref.ref_triangle_ray_intersect(fa(*tri[i]), fa(*to[i]), fa(*td[i])) for i in range(n)], f32)
    # morton, stepping, colour
    xyz = rng.integers(0, 1024, (256, 3)).astype(np.uint32)
    out["morton_xyz"] = xyz
    out["morton"] = np.array([ref.ref_morton3D(C.c_uint32(int(a)), C.c_uint32(int(b)), C.c_uint32(int(c))) for a, b, c in xyz], np.uint32)
    t = np.concatenate([rng.uniform(0, 4, 200), rng.uniform(0, 40, 56)]).astype(f32)
    out["step_t"] = t
    for cone, tag in ((0.0, "c0"), (1.0 / 256.0, "c256")):
        out[f"calc_dt_{tag}"] = np.array([ref.ref_calc_dt(C.c_float(float(x)), C.c_float(cone)) for x in t], f32)
        out[f"adv3_{tag}"] = np.array([ref.ref_advance_n_steps(C.c_float(float(x)), C.c_float(cone), C.c_float(2.5)) for x in t], f32)
    s = rng.uniform(0, 1, 256).astype(f32)
    out["srgb_in"] = s
    out["srgb_lin"] = np.array([ref.ref_srgb_to_linear(C.c_float(float(x))) for x in s], f32)
    dt = rng.uniform(0.0015, 0.2, 256).astype(f32)
    out["dt_in"] = dt
    out["dt_warp"] = np.array([ref.ref_warp_dt(C.c_float(float(x))) for x in dt], f32)
    out["dt_unwarp"] = np.array([ref.ref_unwarp_dt(C.c_float(float(x))) for x in out["dt_warp"]], f32)
    # a-3/a-4/a-6/a-7: pixel -> ray -> start t on the sphere occupancy (bitfield = a-20 on the synthetic grid)
    grid = synthetic.sphere_density_grid(0.35)
    bits, _ = orc.bitfield_from_grid(grid, 1)
    w, h = 48, 40
    cams = {"default": synthetic.default_camera(), "oblique": synthetic.look_at_camera((1.7, 1.2, 1.9), (0.5, 0.5, 0.5))}
    aabb = fa(0, 0, 0, 1, 1, 1)
    for name, cam in cams.items():
        for spp, snap in ((0, 1), (3, 0)):
            p = orc.render_params(w, h, cam, spp=spp, snap=bool(snap))
            ro = np.zeros((h * w, 3), f32); rd = np.zeros((h * w, 3), f32); rt = np.zeros(h * w, f32)
            for y in range(h):
                for x in range(w):
                    oo, dd = (C.c_float * 3)(), (C.c_float * 3)()
                    ref.ref_pixel_ray(C.c_uint32(spp), C.c_uint32(x), C.c_uint32(y), w, h, p.focal, p.cam0, p.cam1, p.rolling_shutter,
                                      p.screen_center, snap, C.c_float(0.0), oo, dd)
                    i = x + w * y
                    ro[i], rd[i] = oo[:], dd[:]
                    rt[i] = ref.ref_start_t(C.c_uint32(spp), C.c_uint32(i), oo, dd, aabb, C.c_float(0.0), P(bits), C.c_uint32(0))
            out[f"ray_o_{name}_{spp}"], out[f"ray_d_{name}_{spp}"], out[f"ray_t_{name}_{spp}"] = ro, rd, rt
    # DDA from random interior starts, 3 cascades, both the unlimited and the 200-iteration (_geometry) variant
    g3 = synthetic.sphere_density_grid(0.6, 3)
    bits3, _ = orc.bitfield_from_grid(g3, 3)
    n = 384
    so = rng.uniform(-1.2, 2.2, (n, 3)).astype(f32)
    sd = (f32(0.5) + rng.normal(scale=0.3, size=(n, 3)).astype(f32) - so).astype(f32)
    sd /= np.linalg.norm(sd, axis=1, keepdims=True).astype(f32)
    st = rng.uniform(0, 0.5, n).astype(f32)
    box3 = fa(-1.5, -1.5, -1.5, 2.5, 2.5, 2.5)
    out["dda_o"], out["dda_d"], out["dda_t0"] = so, sd, st
    for cap in (0, 200):
        for cone, tag in ((0.0, "c0"), (1.0 / 256.0, "c256")):
            out[f"dda_t_{tag}_{cap}"] = np.array([ref.ref_advance_to_occupied(C.c_float(float(st[i])), C.c_float(cone), fa(*so[i]), fa(*sd[i]), P(bits3),
                                                                             C.c_uint32(0), C.c_uint32(2), box3, cap) for i in range(n)], f32)
    np.savez_compressed(os.path.join(HERE, "ref_host.npz"), **out)
    print("ref_host.npz:", len(out), "arrays")


def make_ref_env():
    """a-16 sampling helpers of the reference (random_val.cuh), host-callable -> ref_host_env.npz"""
    ref = orc.ref
    if ref is None or not hasattr(ref, "ref_cosine_hemisphere"):
        raise SystemExit("oracle/_ref/libngp_ref_host.so missing or stale: run `make -C oracle ref`")
    rng = np.random.default_rng(20261105)
    out = {}
    n = 512
    idx = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    seed = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    idx[:16] = np.arange(16)
    v2 = np.zeros((n, 2), f32)
    for i in range(n):
        tmp = (C.c_float * 2)()
        ref.ref_ld_random_val_2d(C.c_uint32(int(idx[i])), C.c_uint32(int(seed[i])), tmp)
        v2[i] = tmp[:]
    out["ld2_idx"], out["ld2_seed"], out["ld2_val"] = idx, seed, v2
    d = rng.normal(size=(n, 3)).astype(f32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(f32)
    d[0] = [0, 0, 1]; d[1] = [0, 0, -1]; d[2] = [1, 0, 0]; d[3] = [0, -1, 0]
    sph = np.zeros((n, 2), f32); frm = np.zeros((n, 9), f32)
    for i in range(n):
        tmp = (C.c_float * 2)(); ref.ref_dir_to_spherical_unorm(fa(*d[i]), tmp); sph[i] = tmp[:]
        t9 = (C.c_float * 9)(); ref.ref_compute_local_frame(fa(*d[i]), t9); frm[i] = t9[:]
    out["dirs"], out["spherical_unorm"], out["local_frame"] = d, sph, frm
    u = rng.uniform(0, 1, (n, 2)).astype(f32)
    u[0] = [0, 0]; u[1] = [1, 1]; u[2] = [0.5, 0.25]
    cos_h = np.zeros((n, 3), f32); cap = np.zeros((n, 3), f32)
    caps = rng.uniform(0.01, 1.0, n).astype(f32)
    for i in range(n):
        t3 = (C.c_float * 3)(); ref.ref_cosine_hemisphere(fa(*u[i]), t3); cos_h[i] = t3[:]
        t3 = (C.c_float * 3)(); ref.ref_spherical_cap(fa(*u[i]), C.c_float(float(caps[i])), t3); cap[i] = t3[:]
    out["u"], out["cosine_hemisphere"], out["cap_height"], out["spherical_cap"] = u, cos_h, caps, cap
    # a-21 loss functions (nerf_device.cuh:74-112,640-655)
    lt_t = rng.random((256, 3)).astype(f32)
    lt_p = (rng.random((256, 3)) * 1.4 - 0.2).astype(f32)
    for lt, tag in ((0, "l2"), (1, "huber"), (2, "relativel2"), (3, "l1"), (4, "mape"), (5, "smape"), (6, "logl1")):
        res = np.zeros((256, 6), f32)
        for i in range(256):
            t6 = (C.c_float * 6)(); ref.ref_loss_and_gradient(fa(*lt_t[i]), fa(*lt_p[i]), lt, t6); res[i] = t6[:]
        out[f"loss_{tag}"] = res
    out["loss_target"], out["loss_pred"] = lt_t, lt_p
    # pcg32 (training ray selection, density-grid jitter): raw outputs after seeding + skip-ahead
    seeds = np.array([0, 1, 1337, 2 ** 40 + 12345], np.uint64)
    advs = np.array([0, 4, 16 * 4096, 2 ** 32], np.uint64)
    pc = np.zeros((4, 4, 8), np.uint32)
    for a, sd in enumerate(seeds):
        for b, ad in enumerate(advs):
            tmp = (C.c_uint32 * 8)()
            ref.ref_pcg32(C.c_uint64(int(sd)), C.c_uint64(int(ad)), 8, tmp)
            pc[a, b] = tmp[:]
    out["pcg_seeds"], out["pcg_advances"], out["pcg_out"] = seeds, advs, pc
    np.savez_compressed(os.path.join(HERE, "ref_host_env.npz"), **out)
    print("ref_host_env.npz:", len(out), "arrays")


LENS_CASES = [  # (tag, lens_mode, lens_params, aperture, focus_z, parallax)
    ("pinhole_dof", 0, (0,) * 7, 0.05, 1.7, (0.02, -0.03, 0.5)),
    ("opencv", 1, (-0.12, 0.05, 0.002, -0.003, 0, 0, 0), 0.0, 1.0, (0, 0, 0)),
    ("ftheta", 2, (0.0, 1.1e-3, 2.0e-8, -1.0e-11, 0.0, 800.0, 600.0), 0.0, 1.0, (0, 0, 0)),
    ("latlong", 3, (0,) * 7, 0.0, 1.0, (0, 0, 0)),
    ("fisheye", 4, (0.03, -0.01, 0.004, -0.001, 0, 0, 0), 0.02, 2.5, (0, 0, 0)),
    ("equirect", 5, (0,) * 7, 0.0, 1.0, (0.01, 0.0, 1.0)),
]


def make_ref_lens():
    """uv_to_ray (common_device.cuh:416-483) of the reference with a Lens, parallax shift and depth of field -> ref_host_lens.npz"""
    ref = orc.ref
    if ref is None or not hasattr(ref, "ref_uv_to_ray_lens"):
        raise SystemExit("oracle/_ref/libngp_ref_host.so missing or stale: run `make -C oracle ref`")
    ref.ref_uv_to_ray_lens.restype = C.c_int32
    rng = np.random.default_rng(20261106)
    n = 192
    uv = rng.uniform(0, 1, (n, 2)).astype(f32)
    uv[0] = [0.5, 0.5]; uv[1] = [0, 0]; uv[2] = [1, 1]; uv[3] = [0.5, 0.0]
    cam = synthetic.look_at_camera((1.4, 1.2, 1.6), (0.5, 0.5, 0.5)).astype(f32).reshape(12)
    w, h = 800, 600
    focal = np.array([700.0, 710.0], f32); sc = np.array([0.48, 0.53], f32)
    out = {"uv": uv, "cam": cam, "res": np.array([w, h], np.int32), "focal": focal, "screen_center": sc, "near": np.array([0.05], f32), "spp": np.array([3], np.uint32)}
    for tag, mode, lp, ap, fz, par in LENS_CASES:
        o = np.zeros((n, 3), f32); d = np.zeros((n, 3), f32); ok = np.zeros(n, np.int32)
        for i in range(n):
            to, td = (C.c_float * 3)(), (C.c_float * 3)()
            ok[i] = ref.ref_uv_to_ray_lens(C.c_uint32(3), fa(*uv[i]), w, h, fa(*focal), fa(*cam), fa(*sc), fa(*par), C.c_float(0.05), C.c_float(fz),
                                           C.c_float(ap), mode, fa(*lp), to, td)
            o[i], d[i] = to[:], td[:]
        out[f"{tag}_o"], out[f"{tag}_d"], out[f"{tag}_valid"] = o, d, ok
    np.savez_compressed(os.path.join(HERE, "ref_host_lens.npz"), **out)
    print("ref_host_lens.npz:", len(out), "arrays")


def make_ref_cdf():
    """binary_search / sample_cdf_2d / image_idx of the reference (common.h:242-264, nerf_device.cuh:538-568,617-639) -> ref_host_cdf.npz.
    The CDFs fed to them are committed in the file (built by the oracle from a seeded error map), so the test needs nothing else."""
    ref = orc.ref
    if ref is None or not hasattr(ref, "ref_sample_cdf_2d"):
        raise SystemExit("oracle/_ref/libngp_ref_host.so missing or stale: run `make -C oracle ref`")
    ref.ref_binary_search.restype = C.c_uint32
    ref.ref_image_idx.restype = C.c_uint32
    rng = np.random.default_rng(20261107)
    n_img, h, w = 3, 12, 16
    data = (rng.random((n_img, h, w)) ** 4).astype(f32)
    data[1, 3:5, 10:13] += 5.0      # a hot spot
    data[2] *= 0.05                 # an image with little error
    data[0, 7] = 0.0                # an empty row
    cx, cy, ci, pi = orc.error_map_cdf(data)
    n = 512
    samples = rng.random((n, 2)).astype(f32)
    samples[:6] = [[0.0, 0.0], [0.5, 0.0], [0.49999997, 0.3], [0.99999994, 0.99999994], [0.75, 0.5], [0.5, 0.99999994]]
    out = {"data": data, "cdf_x_cond_y": cx, "cdf_y": cy, "cdf_img": ci, "samples": samples}
    for img in range(n_img):
        uv = np.zeros((n, 2), f32); pdf = np.zeros(n, f32)
        for i in range(n):
            s = (C.c_float * 2)(*samples[i]); p = C.c_float(0)
            ref.ref_sample_cdf_2d(s, C.c_uint32(img), w, h, P(cx), P(cy), C.byref(p))
            uv[i] = s[:]; pdf[i] = p.value
        out[f"uv_{img}"], out[f"pdf_{img}"] = uv, pdf
    vals = np.concatenate([rng.random(200).astype(f32), ci, np.array([0.0, 1.0, 2.0, -1.0], f32), cy[1][:5]])
    out["bs_vals"] = vals
    out["bs_img"] = np.array([ref.ref_binary_search(C.c_float(v), P(ci), C.c_uint32(n_img)) for v in vals], np.uint32)
    out["bs_row"] = np.array([ref.ref_binary_search(C.c_float(v), P(cy[1]), C.c_uint32(h)) for v in vals], np.uint32)
    base = np.concatenate([np.arange(300), rng.integers(0, 2 ** 31, 100)]).astype(np.uint32)
    out["ii_base"] = base
    for tag, cdf in (("cdf", ci), ("none", None)):
        ii = np.zeros(base.size, np.uint32); pp = np.zeros(base.size, f32)
        for k, b in enumerate(base):
            p = C.c_float(0)
            ii[k] = ref.ref_image_idx(C.c_uint32(int(b)), C.c_uint32(4096), C.c_uint32(n_img), P(cdf) if cdf is not None else None, C.byref(p))
            pp[k] = p.value
        out[f"ii_img_{tag}"], out[f"ii_pdf_{tag}"] = ii, pp
    np.savez_compressed(os.path.join(HERE, "ref_host_cdf.npz"), **out)
    print("ref_host_cdf.npz:", len(out), "arrays")


FOV_CASES = [  # tag, foveation (steepness xy, inverse-piecewise y xy, radius xy), lens mode, lens params
    ("none", (1, 1, 0.5, 0.5, 0.5, 0.5), 0, (0,) * 7),                     # steepness 1 = the fully linear case = Foveation{}
    ("fov_mild", (0.5, 0.6, 0.45, 0.5, 0.2, 0.25), 0, (0,) * 7),
    ("fov_strong", (0.25, 0.3, 0.5, 0.6, 0.15, 0.1), 0, (0,) * 7),
    ("fov_edge", (0.5, 0.5, 0.1, 0.95, 0.3, 0.3), 0, (0,) * 7),          # the linear piece is clipped at 0 / at 1
    ("fov_opencv", (0.5, 0.6, 0.45, 0.5, 0.2, 0.25), 1, (0.05, -0.02, 0.001, -0.002, 0, 0, 0)),
]


def make_ref_fov():
    """Foveation (constructor, warp, unwarp, density), uv_to_ray with foveation + hidden-area mask + distortion grid, pos_to_uv and to_rgb of
    the reference (common_device.cuh:117-246,416-483,505-537,713-739) -> ref_host_fov.npz"""
    ref = orc.ref
    if ref is None or not hasattr(ref, "ref_uv_to_ray_full"):
        raise SystemExit("oracle/_ref/libngp_ref_host.so missing or stale: run `make -C oracle ref`")
    ref.ref_uv_to_ray_full.restype = C.c_int32
    rng = np.random.default_rng(20261108)
    n = 256
    uv = rng.uniform(0, 1, (n, 2)).astype(f32)
    uv[:6] = [[0.5, 0.5], [0, 0], [1, 1], [0.5, 0.0], [0.05, 0.95], [0.999, 0.001]]
    cam = synthetic.look_at_camera((1.4, 1.2, 1.6), (0.5, 0.5, 0.5)).astype(f32).reshape(12)
    w, h = 800, 600
    focal = np.array([700.0, 710.0], f32); sc = np.array([0.48, 0.53], f32); par = np.array([0.01, -0.02, 0.5], f32)
    mask = (rng.random((9, 12)) > 0.25).astype(np.uint8)
    dist = (rng.normal(size=(32, 32, 2)) * 0.02).astype(f32)
    out = {"uv": uv, "cam": cam, "res": np.array([w, h], np.int32), "focal": focal, "screen_center": sc, "parallax": par, "near": np.array([0.05], f32),
           "spp": np.array([3], np.uint32), "mask": mask, "distortion": dist}
    for tag, fov, mode, lp in FOV_CASES:
        co = (C.c_float * 24)()
        ref.ref_foveation_coeffs(fa(*fov), co)
        out[f"{tag}_coeffs"] = np.array(co[:], f32).reshape(2, 12)
        wv, un, de = np.zeros((n, 2), f32), np.zeros((n, 2), f32), np.zeros(n, f32)
        for i in range(n):
            a, b, d = (C.c_float * 2)(), (C.c_float * 2)(), C.c_float(0)
            ref.ref_foveation_warp(fa(*fov), fa(*uv[i]), a, b, C.byref(d))
            wv[i], un[i], de[i] = a[:], b[:], d.value
        out[f"{tag}_warp"], out[f"{tag}_unwarp"], out[f"{tag}_density"] = wv, un, de
        for sub, use_mask, use_dist in (("plain", False, False), ("mask", True, False), ("dist", False, True), ("all", True, True)):
            o = np.zeros((n, 3), f32); d = np.zeros((n, 3), f32); ok = np.zeros(n, np.int32)
            for i in range(n):
                to, td = (C.c_float * 3)(), (C.c_float * 3)()
                ok[i] = ref.ref_uv_to_ray_full(C.c_uint32(3), fa(*uv[i]), w, h, fa(*focal), fa(*cam), fa(*sc), fa(*par), C.c_float(0.05), C.c_float(1.3), C.c_float(0.0),
                                               fa(*fov), P(mask) if use_mask else None, mask.shape[1], mask.shape[0], mode, fa(*lp),
                                               P(dist) if use_dist else None, dist.shape[1], dist.shape[0], to, td)
                o[i], d[i] = to[:], td[:]
            out[f"{tag}_{sub}_o"], out[f"{tag}_{sub}_d"], out[f"{tag}_{sub}_valid"] = o, d, ok
        # pos_to_uv of the point at t = 1 on the un-distorted-grid ray: with a perspective or OpenCV lens this inverts uv_to_ray
        pts = out[f"{tag}_plain_o"] + out[f"{tag}_plain_d"]
        back = np.zeros((n, 2), f32)
        for i in range(n):
            r = (C.c_float * 2)()
            ref.ref_pos_to_uv(fa(*pts[i]), w, h, fa(*focal), fa(*cam), fa(*sc), fa(*par), fa(*fov), mode, fa(*lp), r)
            back[i] = r[:]
        out[f"{tag}_pos_to_uv"] = back
    dirs = (rng.normal(size=(128, 2)) * 0.7).astype(f32)
    dirs[:5] = [[0, 0], [1, 0], [-1, 0], [0, 1], [0, -1e-3]]
    cols = np.zeros((128, 3), f32)
    for i in range(128):
        c = (C.c_float * 3)()
        ref.ref_to_rgb(fa(*dirs[i]), c)
        cols[i] = c[:]
    out["rgb_dirs"], out["rgb"] = dirs, cols
    np.savez_compressed(os.path.join(HERE, "ref_host_fov.npz"), **out)
    print("ref_host_fov.npz:", len(out), "arrays")


def make_ref_tri():
    """Triangle::distance_sq / closest_point (triangle.cuh:66-147) of the reference, host-callable -> ref_host_tri.npz"""
    ref = orc.ref
    if ref is None or not hasattr(ref, "ref_triangle_distance_sq"):
        raise SystemExit("oracle/_ref/libngp_ref_host.so missing or stale: run `make -C oracle ref`")
    ref.ref_triangle_distance_sq.restype = C.c_float
    rng = np.random.default_rng(42)
    n = 512
    tris = rng.uniform(-1, 1, (n, 9)).astype(f32)
    pts = rng.uniform(-1.5, 1.5, (n, 3)).astype(f32)
    for i in range(64):  # points in the triangle's plane, inside it
        w = rng.dirichlet([1, 1, 1]).astype(f32)
        pts[i] = w[0] * tris[i, :3] + w[1] * tris[i, 3:6] + w[2] * tris[i, 6:]
    pts[64] = tris[64, :3]  # a vertex
    d = np.zeros(n, f32); cp = np.zeros((n, 3), f32)
    for i in range(n):
        d[i] = ref.ref_triangle_distance_sq(orc.P(tris[i]), orc.P(pts[i]))
        ref.ref_triangle_closest_point(orc.P(tris[i]), orc.P(pts[i]), orc.P(cp[i]))
    # raystab helpers (random_val.cuh:84-101 fibonacci_dir<32>; default_rng_t advanced by 2 i, triangle_bvh.cu:573-574)
    offs = rng.uniform(0, 1, (8, 2)).astype(f32); offs[0] = 0
    fib = np.zeros((8, 32, 3), f32)
    for a in range(8):
        for i in range(32):
            ref.ref_fibonacci_dir32(C.c_uint32(i), orc.P(offs[a]), orc.P(fib[a, i]))
    adv = np.array([0, 2, 4, 6, 1000, 2 * 4095, 2 ** 33 + 2], np.uint64)
    rng_off = np.zeros((adv.size, 2), f32)
    for a in range(adv.size):
        ref.ref_default_rng_offset(C.c_uint64(int(adv[a])), orc.P(rng_off[a]))
    np.savez_compressed(os.path.join(HERE, "ref_host_tri.npz"), tris=tris, pts=pts, distance_sq=d, closest_point=cp,
                        fib_offsets=offs, fib_dirs=fib, rng_advance=adv, rng_offsets=rng_off)
    print("ref_host_tri.npz: 8 arrays")


class Info:
    pass


def base_info(cfg):
    i = Info()
    i.n_levels, i.n_features_per_level = cfg.n_levels, cfg.n_features_per_level
    i.density_hidden_layers, i.rgb_hidden_layers = cfg.density_hidden_layers, cfg.rgb_hidden_layers
    i.n_params = orc.lib.orc_n_params(C.byref(cfg))
    i.grid_param_offset = orc.lib.orc_grid_param_offset(C.byref(cfg))
    return i


def make_oracle_pin():
    rng = np.random.default_rng(7)
    out = {}
    cfg = orc.net_config(8, 4, 19, 16, float(np.exp(f32(np.log(f32(128.0)) / f32(7.0)))), 1, 2)
    info = base_info(cfg)
    params = synthetic.random_params(info, 1337, 1.0)
    pos = rng.uniform(0, 1, (64, 3)).astype(f32)
    pos[0] = [0, 0, 0]; pos[1] = [1, 1, 1]; pos[2] = [0.5, 0.5, 0.5]
    out["enc_pos"] = pos
    out["enc"] = orc.encode(cfg, params, pos)
    coords = np.concatenate([pos, rng.uniform(0, 1, (64, 1)).astype(f32), rng.uniform(0, 1, (64, 3)).astype(f32)], 1)
    out["net_coords"] = coords
    out["net_out"] = orc.network(cfg, params, coords)
    grid = synthetic.sphere_density_grid(0.35)
    bits, thr = orc.bitfield_from_grid(grid, 1)
    out["bits_sha_like"] = np.array([int(bits.astype(np.uint64).sum()), int((bits.astype(np.uint64) * (np.arange(bits.size) % 8191)).sum())], np.uint64)
    out["bits_thresh"] = np.array([thr], f32)
    tris = synthetic.torus_mesh(48, 24)
    scene = orc.Scene([tris])
    w, h = 40, 30
    p = orc.render_params(w, h, synthetic.default_camera(), mesh_aabb=(-1, -1, -1, 2, 2, 2), n_threads=4)
    frame, depth, steps, total = orc.render_frame(p, cfg, params, bits, scene)
    out["frame"], out["depth"], out["steps"], out["total"] = frame, depth, steps, np.array([total], np.uint64)
    np.savez_compressed(os.path.join(HERE, "oracle_pin.npz"), **out)
    print("oracle_pin.npz:", len(out), "arrays")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "env":  # only the a-16 vectors (leaves the other files untouched)
        make_ref_env()
    elif len(sys.argv) > 1 and sys.argv[1] == "lens":
        make_ref_lens()
    elif len(sys.argv) > 1 and sys.argv[1] == "tri":
        make_ref_tri()
    elif len(sys.argv) > 1 and sys.argv[1] == "cdf":
        make_ref_cdf()
    elif len(sys.argv) > 1 and sys.argv[1] == "fov":
        make_ref_fov()
    else:
        make_ref_fov()
        make_ref_cdf()
        make_ref_tri()
        make_ref_lens()
        make_ref_host()
        make_ref_env()
        make_oracle_pin()
