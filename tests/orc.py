"""ctypes binding of the CPU oracle (oracle/liborc.so) and of the reference's own host functions (oracle/_ref/, when
built). TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu-baseline legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liborc.so")
REF_LIB = os.path.join(ORACLE_DIR, "_ref", "libngp_ref_host.so")

f32, u32, i32 = C.c_float, C.c_uint32, C.c_int32


def _build():
    subprocess.run(["make", "-C", ORACLE_DIR], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)


if not os.path.exists(LIB):
    _build()
lib = C.CDLL(LIB)
ref = C.CDLL(REF_LIB) if os.path.exists(REF_LIB) else None


class NetConfig(C.Structure):
    _fields_ = [("n_levels", u32), ("n_features_per_level", u32), ("log2_hashmap_size", u32), ("base_resolution", u32),
                ("per_level_scale", f32), ("density_hidden_layers", u32), ("rgb_hidden_layers", u32),
                ("has_device_scales", u32), ("device_scales", f32 * 16)]


class RenderParams(C.Structure):
    _fields_ = [("width", i32), ("height", i32), ("spp", u32), ("cam0", f32 * 12), ("cam1", f32 * 12), ("rolling_shutter", f32 * 4),
                ("screen_center", f32 * 2), ("focal", f32 * 2), ("snap_to_pixel_centers", i32), ("near_distance", f32),
                ("render_aabb", f32 * 6), ("train_aabb", f32 * 6), ("max_mip", u32), ("cone_angle_constant", f32),
                ("min_transmittance", f32), ("rgb_activation", i32), ("density_activation", i32), ("linear_colors", i32),
                ("dda_iter_cap", i32), ("mesh_aabb", f32 * 6), ("sun_dir", f32 * 3), ("up_dir", f32 * 3), ("floor_y", f32),
                ("brdf", f32 * 15), ("compat_nearest_aabb_only", i32), ("depth_merge_clip", i32), ("n_threads", i32),
                ("env_w", i32), ("env_h", i32), ("env_samples", i32), ("env_pad", i32), ("envmap", C.c_void_p),
                ("render_mode", i32), ("depth_scale", f32),
                ("lens_mode", i32), ("lens_params", f32 * 7), ("aperture_size", f32), ("focus_z", f32), ("parallax_shift", f32 * 3),
                ("crop_rotated", i32), ("render_aabb_to_local", f32 * 9), ("mask_pad", i32), ("pixel_mask", C.c_void_p),
                ("density_grid_mean", f32), ("pad2", i32),
                ("env_source", i32), ("env_sampling", i32), ("probe_n_theta", i32), ("probe_n_phi", i32), ("probe", C.c_void_p),
                ("foveation", f32 * 6), ("hidden_w", i32), ("hidden_h", i32), ("hidden_area_mask", C.c_void_p),
                ("distortion_w", i32), ("distortion_h", i32), ("distortion", C.c_void_p)]


def P(a):
    return a.ctypes.data_as(C.c_void_p)


for name, res in (("orc_ld_random_val", f32), ("orc_triangle_ray_intersect", f32), ("orc_morton3D", u32), ("orc_advance_to_occupied", f32),
                  ("orc_calc_dt", f32), ("orc_srgb_to_linear", f32), ("orc_bitfield_from_grid", f32), ("orc_n_params", C.c_size_t),
                  ("orc_grid_param_offset", C.c_size_t), ("orc_render_nerf", C.c_uint64), ("orc_render_frame", C.c_uint64),
                  ("orc_scene_create", C.c_void_p)):
    getattr(lib, name).restype = res
lib.orc_scene_destroy.argtypes = [C.c_void_p]
lib.orc_scene_add_mesh.argtypes = [C.c_void_p, C.c_void_p, u32, C.c_void_p]
lib.orc_scene_build.argtypes = [C.c_void_p]
lib.orc_scene_aabb.argtypes = [C.c_void_p, C.c_void_p]
lib.orc_scene_trace.argtypes = [C.c_void_p, C.c_void_p, u32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, i32]
lib.orc_render_mesh.argtypes = [C.c_void_p] * 4
lib.orc_render_frame.argtypes = [C.c_void_p] * 8
lib.orc_render_nerf.argtypes = [C.c_void_p] * 7 + [i32]
lib.orc_init_rays.argtypes = [C.c_void_p] * 6
lib.orc_march_all.argtypes = [C.c_void_p, C.c_void_p, u32] + [C.c_void_p] * 6 + [u32]
lib.orc_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, u32, C.c_void_p]
lib.orc_sh4.argtypes = [C.c_void_p, u32, C.c_void_p]
lib.orc_network.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, u32, C.c_void_p]
lib.orc_bitfield_from_grid.argtypes = [C.c_void_p, u32, C.c_void_p]
lib.orc_n_params.argtypes = [C.c_void_p]
lib.orc_grid_param_offset.argtypes = [C.c_void_p]
lib.orc_grid_layout.argtypes = [C.c_void_p] * 4
lib.orc_ld_random_val_2d.argtypes = [u32, u32, C.c_void_p]
lib.orc_dir_to_spherical_unorm.argtypes = [C.c_void_p, C.c_void_p]
lib.orc_cosine_hemisphere.argtypes = [C.c_void_p, C.c_void_p]
lib.orc_spherical_cap.argtypes = [C.c_void_p, f32, C.c_void_p]
lib.orc_compute_local_frame.argtypes = [C.c_void_p, C.c_void_p]
lib.orc_read_envmap.argtypes = [C.c_void_p, i32, i32, C.c_void_p, C.c_void_p]
lib.orc_env_light_dirs.argtypes = [u32, u32, u32, C.c_void_p, C.c_void_p]
lib.orc_nerf_transmittance.argtypes = [C.c_void_p] * 7
lib.orc_nerf_transmittance.restype = f32
lib.orc_train_gradients.argtypes = [C.c_void_p] * 4 + [u32] + [C.c_void_p] * 5 + [i32, f32, u32, C.c_void_p, C.c_void_p, C.c_void_p]
lib.orc_train_gradients.restype = f32

if ref is not None:
    for name in ("ref_ld_random_val", "ref_triangle_ray_intersect", "ref_advance_to_occupied", "ref_calc_dt", "ref_advance_n_steps",
                 "ref_srgb_to_linear", "ref_warp_dt", "ref_unwarp_dt", "ref_start_t"):
        getattr(ref, name).restype = f32
    ref.ref_morton3D.restype = u32
    if hasattr(ref, "ref_cosine_hemisphere"):
        ref.ref_spherical_cap.argtypes = [C.c_void_p, f32, C.c_void_p]

DEFAULT_BRDF = [0.0, 0.0, 1.0, 0.5, 0.0, 0.0, 0.0, 0.8, 0.8, 0.8, 0.0, 0.0, 0.0, 0.0, 0.0]


def net_config(n_levels=8, n_features=4, log2_T=19, base_res=16, per_level_scale=2.0, density_hidden=1, rgb_hidden=2) -> NetConfig:
    return NetConfig(n_levels, n_features, log2_T, base_res, per_level_scale, density_hidden, rgb_hidden)


def net_config_from_info(info) -> NetConfig:
    """hn_network_info (product) -> oracle config, so both sides use the very same per_level_scale."""
    c = NetConfig(info.n_levels, info.n_features_per_level, info.log2_hashmap_size, info.base_resolution, info.per_level_scale,
                  info.density_hidden_layers, info.rgb_hidden_layers)
    # the per-level scales as the device computes them (MUFU.EX2 + FFMA like the reference's kernel_grid); identical to the
    # host's exp2f for power-of-two per_level_scale
    c.has_device_scales = 1
    for i in range(info.n_levels):
        c.device_scales[i] = info.level_scales[i]
    return c


def with_device_scales(cfg: NetConfig, scales) -> NetConfig:
    """copy of cfg using an explicit per-level scale table (e.g. the one stored beside reference-GPU golden vectors)"""
    c = NetConfig(cfg.n_levels, cfg.n_features_per_level, cfg.log2_hashmap_size, cfg.base_resolution, cfg.per_level_scale,
                  cfg.density_hidden_layers, cfg.rgb_hidden_layers)
    c.has_device_scales = 1
    for i, v in enumerate(scales):
        c.device_scales[i] = float(v)
    return c


def render_params(w, h, cam0, cam1=None, spp=0, rel_focal=None, fov_axis=1, zoom=1.0, screen_center=(0.5, 0.5), snap=True,
                  render_aabb=(0, 0, 0, 1, 1, 1), train_aabb=(0, 0, 0, 1, 1, 1), max_mip=0, cone=0.0, min_transmittance=0.01,
                  rgb_act=2, density_act=3, linear_colors=False, dda_iter_cap=0, mesh_aabb=(0, 0, 0, 1, 1, 1),
                  sun_dir=(0.57735026919, 0.57735026919, 0.57735026919), up_dir=(0, 1, 0), floor_y=-10000.0, brdf=None, compat=False,
                  clip=False, rolling_shutter=(0, 0, 0, 1), near=0.0, n_threads=None, envmap=None, env_samples=0, render_mode=1, depth_scale=1.0 / 0.33,
                  lens_mode=0, lens_params=(0,) * 7, aperture_size=0.0, focus_z=1.0, parallax_shift=(0, 0, 0),
                  render_aabb_to_local=None, pixel_mask=None, density_grid_mean=0.0, probe=None, env_sampling=0,
                  foveation=None, hidden_area_mask=None, distortion=None) -> RenderParams:
    from math import tan, pi
    p = RenderParams()
    p.width, p.height, p.spp = w, h, spp
    cam0 = np.asarray(cam0, dtype=np.float32).reshape(12)
    cam1 = cam0 if cam1 is None else np.asarray(cam1, dtype=np.float32).reshape(12)
    p.cam0[:] = cam0.tolist()
    p.cam1[:] = cam1.tolist()
    p.rolling_shutter[:] = list(rolling_shutter)
    if rel_focal is None:
        rel_focal = np.float32(0.5 / tan(0.5 * 50.625 * pi / 180.0))
        rel_focal = (rel_focal, rel_focal)
    res_axis = np.float32(w if fov_axis == 0 else h)
    # same float32 arithmetic as calc_focal_length / render_screen_center
    p.focal[:] = [float(np.float32(np.float32(np.float32(rel_focal[i]) * res_axis) * np.float32(zoom))) for i in range(2)]
    p.screen_center[:] = [float(np.float32(np.float32(np.float32(0.5) - np.float32(screen_center[i])) * np.float32(zoom)) + np.float32(0.5)) for i in range(2)]
    p.snap_to_pixel_centers = int(snap)
    p.near_distance = near
    p.render_aabb[:] = list(render_aabb)
    p.train_aabb[:] = list(train_aabb)
    p.max_mip, p.cone_angle_constant, p.min_transmittance = max_mip, cone, min_transmittance
    p.rgb_activation, p.density_activation, p.linear_colors, p.dda_iter_cap = rgb_act, density_act, int(linear_colors), dda_iter_cap
    p.mesh_aabb[:] = list(mesh_aabb)
    p.sun_dir[:] = list(sun_dir)
    p.up_dir[:] = list(up_dir)
    p.floor_y = floor_y
    p.brdf[:] = list(brdf) if brdf is not None else DEFAULT_BRDF
    p.compat_nearest_aabb_only, p.depth_merge_clip = int(compat), int(clip)
    p.n_threads = n_threads if n_threads is not None else (os.cpu_count() or 1)
    p.render_mode, p.depth_scale = int(render_mode), float(depth_scale)
    p.lens_mode, p.aperture_size, p.focus_z = int(lens_mode), float(aperture_size), float(focus_z)
    p.lens_params[:] = [float(v) for v in lens_params]
    p.parallax_shift[:] = [float(v) for v in parallax_shift]
    ident = [1, 0, 0, 0, 1, 0, 0, 0, 1]
    R = ident if render_aabb_to_local is None else [float(v) for v in np.asarray(render_aabb_to_local, np.float32).reshape(9)]
    p.render_aabb_to_local[:] = R
    p.crop_rotated = int(any(abs(a - b) > 0 for a, b in zip(R, ident)))
    p.density_grid_mean = float(density_grid_mean)
    if foveation is not None:       # center_pixel_steepness xy, center_inverse_piecewise_y xy, center_radius xy
        p.foveation[:] = [float(v) for v in foveation]
    if hidden_area_mask is not None:   # uint8 [h][w], 0 = hidden
        hm = np.ascontiguousarray(hidden_area_mask, np.uint8)
        p._hidden_keep = hm
        p.hidden_area_mask, p.hidden_h, p.hidden_w = hm.ctypes.data, hm.shape[0], hm.shape[1]
    if distortion is not None:         # float32 [h][w][2]
        dm = np.ascontiguousarray(distortion, np.float32)
        assert dm.ndim == 3 and dm.shape[2] == 2
        p._dist_keep = dm
        p.distortion, p.distortion_h, p.distortion_w = dm.ctypes.data, dm.shape[0], dm.shape[1]
    if probe is not None:  # a-16: the NeRF's probe texture [n_phi][n_theta][4] lights the meshes
        pt = np.ascontiguousarray(probe, np.float32)
        assert pt.ndim == 3 and pt.shape[2] == 4
        p._probe_keep = pt
        p.probe = pt.ctypes.data
        p.probe_n_phi, p.probe_n_theta = pt.shape[0], pt.shape[1]
        p.env_source, p.env_sampling, p.env_samples = 1, int(env_sampling), int(env_samples)
    if pixel_mask is not None:
        pm = np.ascontiguousarray(pixel_mask, np.uint8).reshape(-1)
        assert pm.size == w * h
        p._mask_keep = pm
        p.pixel_mask = pm.ctypes.data
    if envmap is not None:
        env = np.ascontiguousarray(envmap, np.float32)
        assert env.ndim == 3 and env.shape[2] == 4
        p._env_keep = env  # keep the buffer alive as long as the params
        p.env_h, p.env_w, p.env_samples = env.shape[0], env.shape[1], int(env_samples)
        p.envmap = env.ctypes.data
    return p


class Scene:
    def __init__(self, meshes):
        self.h = C.c_void_p(lib.orc_scene_create())
        self._keep = []
        for t in meshes:
            t = np.ascontiguousarray(t, dtype=np.float32).reshape(-1, 9)
            c = np.zeros(3, np.float32)
            self._keep.append(t)
            lib.orc_scene_add_mesh(self.h, P(t), t.shape[0], P(c))
        lib.orc_scene_build(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib.orc_scene_destroy(self.h)
            self.h = None

    def aabb(self):
        out = np.zeros(6, np.float32)
        lib.orc_scene_aabb(self.h, P(out))
        return out

    def distance(self, positions, mode=0, max_distance=10.0, n_threads=None):
        """brute-force unsigned (mode 0) / watertight-signed (mode 1) distance of every position to the scene; (distance, triangle)"""
        pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        out = np.empty(pos.shape[0], np.float32); tri = np.empty(pos.shape[0], np.int32)
        lib.orc_scene_distance(self.h, C.c_uint32(pos.shape[0]), P(pos), C.c_int32(mode), C.c_float(max_distance), C.c_int32(n_threads or (os.cpu_count() or 1)), P(out), P(tri))
        return out, tri

    def trace(self, p: RenderParams, origins, dirs, brute=False):
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        n = o.shape[0]
        mesh = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
        lib.orc_scene_trace(self.h, C.byref(p), n, P(o), P(d), P(mesh), P(tri), P(t), int(brute))
        return mesh, tri, t


def bitfield_from_grid(grid, n_cascades=1):
    g = np.ascontiguousarray(grid, np.float32)
    out = np.zeros(128 ** 3 // 8 * 8, np.uint8)
    thresh = lib.orc_bitfield_from_grid(P(g), n_cascades, P(out))
    return out, thresh


def init_rays(p: RenderParams, bits):
    n = p.width * p.height
    o = np.empty((n, 3), np.float32); d = np.empty((n, 3), np.float32); t = np.empty(n, np.float32); a = np.empty(n, np.uint8)
    lib.orc_init_rays(C.byref(p), P(bits), P(o), P(d), P(t), P(a))
    return o, d, t, a


def march_all(p: RenderParams, bits, o, d, t0, alive, max_per_ray=0):
    n = t0.size
    counts = np.empty(n, np.uint32)
    samples = np.zeros((n, max_per_ray, 4), np.float32) if max_per_ray else None
    lib.orc_march_all(C.byref(p), P(bits), n, P(np.ascontiguousarray(o, np.float32)), P(np.ascontiguousarray(d, np.float32)),
                      P(np.ascontiguousarray(t0, np.float32)), P(np.ascontiguousarray(alive, np.uint8)), P(counts),
                      P(samples) if max_per_ray else None, max_per_ray)
    return counts, samples


def encode(cfg: NetConfig, params, positions):
    pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
    out = np.empty((pos.shape[0], 32), np.float16)
    lib.orc_encode(C.byref(cfg), P(params), P(pos), pos.shape[0], P(out))
    return out


def sh4(dirs01):
    """degree-4 SH of warped directions in [0,1]^3 -> fp16 [n][16]"""
    d = np.ascontiguousarray(dirs01, np.float32).reshape(-1, 3)
    out = np.empty((d.shape[0], 16), np.float16)
    lib.orc_sh4(P(d), C.c_uint32(d.shape[0]), P(out))
    return out


def normalize_mesh(tris, center=(0.0, 0.0, 0.0)):
    """Testbed::load_mesh's normalisation (src/testbed_geometry_training.cpp:2805-2823) in float32, same operation order:
    box of all vertices inflated by 0.5 % of its diagonal, longest side -> 1, centred on 0.5, then offset by `center`."""
    f = np.float32
    v = np.ascontiguousarray(tris, np.float32).reshape(-1, 3)
    lo, hi = v.min(0), v.max(0)
    diag = hi - lo
    ln = f(0.0)
    for d in range(3):
        ln = f(ln + f(diag[d] * diag[d]))
    infl = f(np.sqrt(ln) * f(0.005))
    lo = (lo - infl).astype(f); hi = (hi + infl).astype(f)
    diag = (hi - lo).astype(f)
    scale = f(diag.max())
    x = ((((v - lo).astype(f) - (f(0.5) * diag).astype(f)).astype(f) / scale).astype(f) + f(0.5)).astype(f)
    x = (x + np.asarray(center, f)).astype(f)
    return np.ascontiguousarray(x.reshape(-1, 9))


def network(cfg: NetConfig, params, coords):
    c = np.ascontiguousarray(coords, np.float32).reshape(-1, 7)
    out = np.empty((c.shape[0], 4), np.float16)
    lib.orc_network(C.byref(cfg), P(params), P(c), c.shape[0], P(out))
    return out


def read_envmap(env, dirs):
    env = np.ascontiguousarray(env, np.float32)
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    out = np.empty((d.shape[0], 4), np.float32)
    for i in range(d.shape[0]):
        lib.orc_read_envmap(P(env), env.shape[1], env.shape[0], P(d[i]), P(out[i]))
    return out


def env_light_dirs(idx, spp, K, normal):
    n = np.ascontiguousarray(normal, np.float32)
    out = np.empty((K, 3), np.float32)
    lib.orc_env_light_dirs(int(idx), int(spp), int(K), P(n), P(out))
    return out


def nerf_transmittance(p: RenderParams, cfg, params, bits, o, d):
    o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
    n = C.c_uint32(0)
    T = lib.orc_nerf_transmittance(C.byref(p), C.byref(cfg), P(params), P(bits), P(o), P(d), C.byref(n))
    return float(T), int(n.value)


def probe_rays(mode, n_theta, n_phi, n_origin=1, box=(0, 0, 0, 1, 1, 1), origin=(0.0, 0.0, 0.0)):
    n = n_theta * n_phi * (n_origin * n_origin if mode == 2 else 1)
    o = np.zeros((n, 3), np.float32); d = np.zeros((n, 3), np.float32); texel = np.zeros(n, np.uint32)
    b = np.asarray(box, np.float32); og = np.asarray(origin, np.float32)
    lib.orc_probe_rays(C.c_int32(mode), u32(n_theta), u32(n_phi), u32(n_origin), P(b), P(og), P(o), P(d), P(texel))
    return o, d, texel


def render_probe(p: RenderParams, cfg, params, bits, mode, n_theta, n_phi, n_origin=1, origin=(0.0, 0.0, 0.0)):
    """oracle probe texture [n_phi][n_theta][4] (premultiplied linear RGBA) and the number of network samples"""
    tex = np.zeros((n_phi, n_theta, 4), np.float32)
    og = np.asarray(origin, np.float32)
    lib.orc_render_probe.restype = C.c_uint64
    n = lib.orc_render_probe(C.byref(p), C.byref(cfg), P(params), P(bits), C.c_int32(mode), u32(n_theta), u32(n_phi), u32(n_origin), P(og), P(tex))
    return tex, int(n)


def read_probe(tex, dirs):
    t = np.ascontiguousarray(tex, np.float32)
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    out = np.zeros((d.shape[0], 4), np.float32)
    lib.orc_read_probe(P(t), u32(t.shape[1]), u32(t.shape[0]), P(d), u32(d.shape[0]), P(out))
    return out


def halton(base, idx):
    lib.orc_halton.restype = C.c_float
    return float(lib.orc_halton(u32(base), C.c_uint64(idx)))


def cylindrical_to_dir(p2):
    a = np.asarray(p2, np.float32); out = np.zeros(3, np.float32)
    lib.orc_cylindrical_to_dir(P(a), P(out))
    return out


def train_gradients(p: RenderParams, cfg, params, bits, origins, dirs, t_start, targets, backgrounds, loss_type=1, loss_scale=128.0, max_samples=1 << 20):
    """a-21 oracle: (mean loss, fp32 gradients [n_params] scaled by loss_scale, samples used per ray)"""
    o = np.ascontiguousarray(origins, np.float32); d = np.ascontiguousarray(dirs, np.float32); t = np.ascontiguousarray(t_start, np.float32)
    tg = np.ascontiguousarray(targets, np.float32); bg = np.ascontiguousarray(backgrounds, np.float32)
    n = o.shape[0]
    grads = np.zeros(int(lib.orc_n_params(C.byref(cfg))), np.float32)
    steps = np.zeros(n, np.uint32)
    total = u32(0)
    loss = lib.orc_train_gradients(C.byref(p), C.byref(cfg), P(params), P(bits), n, P(o), P(d), P(t), P(tg), P(bg), int(loss_type), float(loss_scale),
                                   int(max_samples), P(grads), P(steps), C.byref(total))
    return float(loss), grads, steps


def train_gradients_depth(p: RenderParams, cfg, params, bits, origins, dirs, t_start, targets, backgrounds, target_depth, depth_lambda, depth_loss_type=3,
                          loss_type=1, loss_scale=128.0, max_samples=1 << 20):
    """train_gradients with depth supervision: target_depth [n] (<= 0: none), lambda, ELossType of the depth term (3 = L1, the reference's default)"""
    o = np.ascontiguousarray(origins, np.float32); d = np.ascontiguousarray(dirs, np.float32); t = np.ascontiguousarray(t_start, np.float32)
    tg = np.ascontiguousarray(targets, np.float32); bg = np.ascontiguousarray(backgrounds, np.float32); td = np.ascontiguousarray(target_depth, np.float32)
    n = o.shape[0]
    grads = np.zeros(int(lib.orc_n_params(C.byref(cfg))), np.float32)
    steps = np.zeros(n, np.uint32)
    total = u32(0)
    lib.orc_train_gradients_depth.argtypes = [C.c_void_p] * 4 + [u32] + [C.c_void_p] * 6 + [f32, i32, i32, f32, u32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orc_train_gradients_depth.restype = f32
    loss = lib.orc_train_gradients_depth(C.byref(p), C.byref(cfg), P(params), P(bits), n, P(o), P(d), P(t), P(tg), P(bg), P(td), float(depth_lambda), int(depth_loss_type),
                                         int(loss_type), float(loss_scale), int(max_samples), P(grads), P(steps), C.byref(total))
    return float(loss), grads, steps


def update_density_grid(p: RenderParams, cfg, params, grid, n_cascades, ema_step, n_uniform, n_nonuniform, decay=0.95, seed=1337, rng_calls=0):
    g = np.ascontiguousarray(grid, np.float32).copy()
    lib.orc_update_density_grid.argtypes = [C.c_void_p] * 4 + [u32, u32, u32, u32, f32, u32, u32]
    lib.orc_update_density_grid(C.byref(p), C.byref(cfg), P(params), P(g), n_cascades, ema_step, n_uniform, n_nonuniform, decay, seed, rng_calls)
    return g


def mark_untrained_grid(n_cascades, cams, w, h, focal, principal=(0.5, 0.5)):
    g = np.zeros(128 ** 3 * n_cascades, np.float32)
    c = np.ascontiguousarray(cams, np.float32).reshape(-1, 12)
    f = np.asarray([focal, focal] if np.isscalar(focal) else focal, np.float32)
    pp = np.asarray(principal, np.float32)
    lib.orc_mark_untrained_grid.argtypes = [C.c_void_p, u32, u32, C.c_void_p, i32, i32, C.c_void_p, C.c_void_p]
    lib.orc_mark_untrained_grid(P(g), n_cascades, c.shape[0], P(c), w, h, P(f), P(pp))
    return g


def mark_untrained_grid_per_image(n_cascades, cams, w, h, focal, principal, lens_mode=None, lens_params=None):
    g = np.zeros(128 ** 3 * n_cascades, np.float32)
    c = np.ascontiguousarray(cams, np.float32).reshape(-1, 12)
    f, pp = np.ascontiguousarray(focal, np.float32).reshape(-1, 2), np.ascontiguousarray(principal, np.float32).reshape(-1, 2)
    lm = np.ascontiguousarray(lens_mode, np.int32) if lens_mode is not None else None
    lp = np.ascontiguousarray(lens_params, np.float32).reshape(-1, 7) if lens_params is not None else None
    lib.orc_mark_untrained_grid_per_image.argtypes = [C.c_void_p, u32, u32, C.c_void_p, i32, i32] + [C.c_void_p] * 4
    lib.orc_mark_untrained_grid_per_image(P(g), n_cascades, c.shape[0], P(c), w, h, P(f), P(pp), P(lm) if lm is not None else None, P(lp) if lp is not None else None)
    return g


def render_frame(p: RenderParams, cfg, params, bits, scene: Scene | None):
    n = p.width * p.height
    frame = np.zeros((p.height, p.width, 4), np.float32)
    depth = np.zeros((p.height, p.width), np.float32)
    steps = np.zeros(n, np.uint32)
    total = lib.orc_render_frame(scene.h if scene is not None else None, C.byref(p), C.byref(cfg) if cfg is not None else None,
                                 P(params) if params is not None else None, P(bits) if bits is not None else None, P(frame), P(depth), P(steps))
    return frame, depth, steps.reshape(p.height, p.width), int(total)


def tonemap(accum, to_srgb, exposure, bg, curve, clamp_output=True, unmultiply_alpha=False):
    """oracle tonemap_kernel on an (..., 4) linear accumulation buffer"""
    a = np.ascontiguousarray(accum, dtype=np.float32)
    out = np.empty_like(a)
    bgv = np.asarray(bg, np.float32)
    lib.orc_tonemap(P(a), C.c_uint32(a.size // 4), C.c_int32(int(to_srgb)), C.c_float(float(exposure)), P(bgv), C.c_int32(int(curve)),
                    C.c_int32(int(clamp_output)), C.c_int32(int(unmultiply_alpha)), P(out))
    return out


def uv_to_ray(p: RenderParams, spp, uv, use_lens=True):
    """oracle uv_to_ray for an array of uv (camera = p.cam0): (origins, directions (not normalised), valid)"""
    uv = np.ascontiguousarray(uv, dtype=np.float32)
    n = uv.shape[0]
    o = np.zeros((n, 3), np.float32); d = np.zeros((n, 3), np.float32); ok = np.zeros(n, np.int32)
    lib.orc_uv_to_ray.restype = C.c_int32
    for i in range(n):
        ok[i] = lib.orc_uv_to_ray(C.byref(p), C.c_uint32(int(spp)), P(uv[i]), C.c_int32(int(use_lens)), P(o[i]), P(d[i]))
    return o, d, ok


def render_reference_frame(tb, params, grid, n_cascades, meshes, cam, w, h, mesh_aabb=None, spp=0, compat=False, clip=False,
                           n_threads=None, **kw):
    """Oracle frame with the same settings as a product Testbed `tb` (network info, options) for parity checks.
    Returns (frame, depth, n_steps) in the product's LINEAR accumulate space (spp = 1, no tonemap)."""
    cfg = net_config_from_info(tb.network_info()) if params is not None else None
    bits = bitfield_from_grid(grid, n_cascades)[0] if grid is not None else None
    scene = Scene(meshes) if meshes else None
    if mesh_aabb is None and scene is not None:
        bb = scene.aabb()
        mesh_aabb = np.concatenate([bb[:3] - np.float32(4.0), bb[3:] + np.float32(4.0)])
    p = render_params(w, h, cam, spp=spp, snap=bool(tb.get_option("snap_to_pixel_centers")),
                      rel_focal=tb.get_option("relative_focal_length", 2), fov_axis=int(tb.get_option("fov_axis")), zoom=tb.get_option("zoom"),
                      render_aabb=tb.get_option("render_aabb", 6), train_aabb=tb.get_option("nerf_aabb", 6),
                      max_mip=int(tb.get_option("max_cascade")), cone=tb.get_option("cone_angle_constant"),
                      min_transmittance=tb.get_option("render_min_transmittance"), linear_colors=bool(tb.get_option("linear_colors")),
                      mesh_aabb=mesh_aabb if mesh_aabb is not None else (0, 0, 0, 1, 1, 1), compat=compat, clip=clip, n_threads=n_threads, **kw)
    frame, depth, steps, _ = render_frame(p, cfg, params, bits, scene)
    # product epilogue for spp=1, linear: background (0,0,0,1) blend adds alpha only, then clamp
    out = frame.copy()
    out[..., 3] = out[..., 3] + (1.0 - out[..., 3]) * 1.0
    out = np.clip(out, 0.0, 1.0)
    return out, depth, steps


# ---- training error map (oracle/orc_errmap.inc)
def binary_search(val, data):
    d = np.ascontiguousarray(data, np.float32)
    lib.orc_binary_search.restype = C.c_uint32
    return int(lib.orc_binary_search(C.c_float(val), P(d), u32(d.size)))


def image_idx(base_idx, n_rays, n_images, cdf=None):
    pdf = C.c_float(0)
    c = np.ascontiguousarray(cdf, np.float32) if cdf is not None else None
    lib.orc_image_idx.restype = C.c_uint32
    img = lib.orc_image_idx(u32(base_idx), u32(n_rays), u32(n_images), P(c) if c is not None else None, C.byref(pdf))
    return int(img), float(pdf.value)


def sample_cdf_2d(sample, img, cdf_x_cond_y, cdf_y):
    cx, cy = np.ascontiguousarray(cdf_x_cond_y, np.float32), np.ascontiguousarray(cdf_y, np.float32)
    s = np.array(sample, np.float32); pdf = C.c_float(0)
    lib.orc_sample_cdf_2d(P(s), u32(img), C.c_int32(cx.shape[2]), C.c_int32(cx.shape[1]), P(cx), P(cy), C.byref(pdf))
    return s, float(pdf.value)


def error_map_cdf(data):
    d = np.ascontiguousarray(data, np.float32)
    n, h, w = d.shape
    cx, cy, ci, pi = np.empty_like(d), np.empty((n, h), np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
    lib.orc_error_map_cdf(u32(n), u32(h), u32(w), P(d), P(cx), P(cy), P(ci), P(pi))
    return cx, cy, ci, pi


def sample_training_pixels(n_rays, n_images, seed=1337, step=0, ray_offset=0, n_rays_total=None, cdf_x_cond_y=None, cdf_y=None, cdf_img=None):
    total = n_rays if n_rays_total is None else n_rays_total
    cx = np.ascontiguousarray(cdf_x_cond_y, np.float32) if cdf_x_cond_y is not None else None
    cy = np.ascontiguousarray(cdf_y, np.float32) if cdf_y is not None else None
    ci = np.ascontiguousarray(cdf_img, np.float32) if cdf_img is not None else None
    rx, ry = (cx.shape[2], cx.shape[1]) if cx is not None else (0, 0)
    img, uv, pdf = np.empty(n_rays, np.uint32), np.empty((n_rays, 2), np.float32), np.empty(n_rays, np.float32)
    lib.orc_sample_training_pixels(u32(n_rays), u32(ray_offset), u32(total), u32(seed), u32(step), u32(n_images), C.c_int32(rx), C.c_int32(ry),
                                   P(cx) if cx is not None else None, P(cy) if cy is not None else None, P(ci) if ci is not None else None, P(img), P(uv), P(pdf))
    return img, uv, pdf


def error_map_deposit(img, uv, val, n_images, res, image_res):
    i, u, v = np.ascontiguousarray(img, np.uint32).reshape(-1), np.ascontiguousarray(uv, np.float32).reshape(-1, 2), np.ascontiguousarray(val, np.float32).reshape(-1)
    out = np.zeros((n_images, res[1], res[0]), np.float32)
    lib.orc_error_map_deposit(u32(i.size), P(i), P(u), P(v), C.c_int32(res[0]), C.c_int32(res[1]), C.c_int32(image_res[0]), C.c_int32(image_res[1]), P(out))
    return out


# ---- foveation / 2-D helpers (oracle/orc_fov.inc)
def foveation_coeffs(f6):
    f = np.asarray(f6, np.float32); out = np.zeros(24, np.float32)
    lib.orc_foveation_coeffs(P(f), P(out))
    return out.reshape(2, 12)


def foveation_warp(f6, uv):
    f = np.asarray(f6, np.float32); u = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
    w, un, dens = np.zeros_like(u), np.zeros_like(u), np.zeros(u.shape[0], np.float32)
    for i in range(u.shape[0]):
        d = C.c_float(0)
        lib.orc_foveation_warp(P(f), P(u[i]), P(w[i]), P(un[i]), C.byref(d))
        dens[i] = d.value
    return w, un, dens


def pos_to_uv(p: RenderParams, pos):
    q = np.ascontiguousarray(pos, np.float32).reshape(-1, 3); out = np.zeros((q.shape[0], 2), np.float32)
    for i in range(q.shape[0]):
        lib.orc_pos_to_uv(C.byref(p), P(q[i]), P(out[i]))
    return out


def to_rgb(dirs):
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 2); out = np.zeros((d.shape[0], 3), np.float32)
    for i in range(d.shape[0]):
        lib.orc_to_rgb(P(d[i]), P(out[i]))
    return out
