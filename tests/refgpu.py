"""ctypes binding of oracle/_ref/libngp_ref_gpu.so — the reference's OWN CUDA build (ngp::Testbed + tiny-cuda-nn compiled
for sm_100a by oracle/ref_gpu.mk, C wrapper oracle/ref_gpu_shim.cu). TEST / MEASUREMENT INFRASTRUCTURE ONLY: used by
tests/golden/make_ref_gpu_golden.py (golden vectors), by the live cross-checks in tests/test_gpu_refgpu.py and by
bench.py's reference arm. The product never loads it."""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_ref", "libngp_ref_gpu.so")

_lib = None


def available() -> bool:
    """True when the reference CUDA build exists AND loads (it needs the driver's libcuda.so.1, i.e. a GPU box)."""
    if not os.path.exists(LIB_PATH):
        return False
    try:
        lib()
        return True
    except OSError:
        return False


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError(f"{LIB_PATH} missing: build it with `make -C oracle -f ref_gpu.mk -j8` where /root/reference is mounted")
        _lib = C.CDLL(LIB_PATH)
        _lib.rg_last_error.restype = C.c_char_p
        _lib.rg_create.restype = C.c_void_p
        _lib.rg_destroy.argtypes = [C.c_void_p]
        for n in ("rg_load_training_data", "rg_load_snapshot", "rg_reload_network"):
            getattr(_lib, n).argtypes = [C.c_void_p, C.c_char_p]
        _lib.rg_save_snapshot.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        _lib.rg_set_mode.argtypes = [C.c_void_p, C.c_int]
        _lib.rg_mode.argtypes = [C.c_void_p]
        _lib.rg_set_camera.argtypes = [C.c_void_p, C.c_void_p]
        _lib.rg_set.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int]
        _lib.rg_get.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int]
        _lib.rg_get_bitfield.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        _lib.rg_get_params.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        _lib.rg_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.rg_time_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_void_p]
        _lib.rg_network_inference.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        _lib.rg_encode_pos.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        _lib.rg_encode_dir.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        _lib.rg_encoding_widths.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.rg_time_inference.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_void_p]
        _lib.rg_trace_mesh.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        _lib.rg_mesh_triangles.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_size_t]
        _lib.rg_mesh_count.argtypes = [C.c_void_p]
        _lib.rg_mesh_count.restype = C.c_uint32
        _lib.rg_mesh_n_triangles.argtypes = [C.c_void_p, C.c_uint32]
        _lib.rg_mesh_n_triangles.restype = C.c_uint64
        _lib.rg_probe.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 7
        _lib.rg_train.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
        _lib.rg_time_train.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


BATCH = 256  # tcnn BATCH_SIZE_GRANULARITY


def pad_rows(a: np.ndarray, mult: int = BATCH) -> np.ndarray:
    n = a.shape[0]
    m = (n + mult - 1) // mult * mult
    if m == n:
        return a
    return np.concatenate([a, np.repeat(a[-1:], m - n, axis=0)], 0)


class RefTestbed:
    """The reference's ngp::Testbed, headless."""

    def __init__(self):
        self.l = lib()
        self.h = self.l.rg_create()
        if not self.h:
            raise RuntimeError("rg_create: " + self.l.rg_last_error().decode())

    def close(self):
        if getattr(self, "h", None):
            self.l.rg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise RuntimeError(self.l.rg_last_error().decode())

    def load_training_data(self, path):
        self._ck(self.l.rg_load_training_data(self.h, os.fsencode(path)))

    def load_snapshot(self, path):
        self._ck(self.l.rg_load_snapshot(self.h, os.fsencode(path)))

    def save_snapshot(self, path, with_optimizer=False):
        self._ck(self.l.rg_save_snapshot(self.h, os.fsencode(path), int(with_optimizer)))

    def reload_network(self, path):
        self._ck(self.l.rg_reload_network(self.h, os.fsencode(path)))

    def set_mode(self, mode: int):
        self._ck(self.l.rg_set_mode(self.h, mode))

    def mode(self) -> int:
        return int(self.l.rg_mode(self.h))

    def set_camera(self, cols12):
        c = _f32(cols12).reshape(12)
        self._ck(self.l.rg_set_camera(self.h, _p(c)))

    def set(self, name, value):
        v = _f32(np.atleast_1d(value))
        self._ck(self.l.rg_set(self.h, name.encode(), _p(v), v.size))

    def get(self, name, n=1):
        v = np.zeros(max(n, 1), np.float32)
        self._ck(self.l.rg_get(self.h, name.encode(), _p(v), n))
        return v if n > 1 else float(v[0])

    def bitfield(self, n_bytes=128 ** 3 // 8 * 8):
        out = np.zeros(n_bytes, np.uint8)
        self._ck(self.l.rg_get_bitfield(self.h, _p(out), out.size))
        return out

    def params(self, n):
        out = np.zeros(n, np.float16)
        self._ck(self.l.rg_get_params(self.h, _p(out), n))
        return out

    def render(self, w, h, spp=1, linear=True, raw=False):
        """-> tonemapped RGBA [h,w,4] (what pyngp.Testbed.render returns); with raw=True also the raw frame buffer
        (premultiplied linear RGBA of the last sample) and the depth buffer."""
        out = np.zeros((h, w, 4), np.float32)
        fr = np.zeros((h, w, 4), np.float32) if raw else None
        dp = np.zeros((h, w), np.float32) if raw else None
        self._ck(self.l.rg_render(self.h, w, h, spp, int(linear), _p(out), _p(fr), _p(dp)))
        return (out, fr, dp) if raw else out

    def time_render(self, w, h, warmup=3, steps=20, l2_flush_bytes=512 << 20):
        ms = np.zeros(steps, np.float32)
        self._ck(self.l.rg_time_render(self.h, w, h, warmup, steps, l2_flush_bytes, _p(ms)))
        return ms

    def widths(self):
        a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._ck(self.l.rg_encoding_widths(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def network(self, coords7):
        """coords7: [n,7] (warped pos, dt, warped dir) -> raw fp16 (r,g,b,sigma) [n,4]."""
        c = _f32(coords7).reshape(-1, 7)
        n = c.shape[0]
        cp = pad_rows(c)
        out = np.zeros((4, cp.shape[0]), np.float16)
        self._ck(self.l.rg_network_inference(self.h, _p(cp), cp.shape[0], _p(out)))
        return np.ascontiguousarray(out[:, :n].T)

    def encode_pos(self, pos):
        p = _f32(pos).reshape(-1, 3)
        n = p.shape[0]
        pp = pad_rows(p)
        wdt = self.widths()[0]
        out = np.zeros((wdt, pp.shape[0]), np.float16)
        self._ck(self.l.rg_encode_pos(self.h, _p(pp), pp.shape[0], _p(out)))
        return np.ascontiguousarray(out[:, :n].T)

    def encode_dir(self, dirs01):
        d = _f32(dirs01).reshape(-1, 3)
        n = d.shape[0]
        dp = pad_rows(d)
        wdt = self.widths()[1]
        out = np.zeros((wdt, dp.shape[0]), np.float16)
        self._ck(self.l.rg_encode_dir(self.h, _p(dp), dp.shape[0], _p(out)))
        return np.ascontiguousarray(out[:, :n].T)

    def time_inference(self, coords_dev_ptr: int, n_total: int, chunk: int = 2 << 20, reps: int = 3) -> float:
        ms = C.c_float(0)
        self._ck(self.l.rg_time_inference(self.h, C.c_void_p(coords_dev_ptr), n_total, chunk, reps, C.byref(ms)))
        return float(ms.value)

    def trace_mesh(self, origins, dirs):
        """mesh_raytrace_kernel: -> (positions after the trace, directions/normals after the trace)."""
        o = _f32(origins).reshape(-1, 3).copy()
        d = _f32(dirs).reshape(-1, 3).copy()
        self._ck(self.l.rg_trace_mesh(self.h, _p(o), _p(d), o.shape[0]))
        return o, d

    def mesh_triangles(self, m):
        n = int(self.l.rg_mesh_n_triangles(self.h, m))
        out = np.zeros((n, 9), np.float32)
        self._ck(self.l.rg_mesh_triangles(self.h, m, _p(out), n))
        return out

    def mesh_count(self):
        return int(self.l.rg_mesh_count(self.h))

    def probe(self, mode, n_theta, n_phi, n_origin=1, origin=(0.0, 0.0, 0.0)):
        """the reference's probe path: init_rays_from_center{,_outward,_multiple_center} + NerfTracer::trace_mesh.
        -> (ray origins, ray dirs, ray texel idx, hit texel idx, hit rgba)"""
        n = n_theta * n_phi * (n_origin * n_origin if mode == 2 else 1)
        o = np.zeros((n, 3), np.float32); d = np.zeros((n, 3), np.float32); idx = np.zeros(n, np.uint32)
        hidx = np.zeros(n, np.uint32); hrgba = np.zeros((n, 4), np.float32)
        nh = C.c_uint32(0)
        og = _f32(origin).reshape(3)
        self._ck(self.l.rg_probe(self.h, int(mode), n_theta, n_phi, n_origin, _p(og), _p(o), _p(d), _p(idx), C.byref(nh), _p(hidx), _p(hrgba)))
        return o, d, idx, hidx[: nh.value], hrgba[: nh.value]

    def train(self, batch=1 << 18, n_steps=1):
        loss = C.c_float(0)
        self._ck(self.l.rg_train(self.h, batch, n_steps, C.byref(loss)))
        return float(loss.value)

    def time_train(self, batch=1 << 18, warmup=8, steps=20):
        ms, sm = C.c_float(0), C.c_float(0)
        self._ck(self.l.rg_time_train(self.h, batch, warmup, steps, C.byref(ms), C.byref(sm)))
        return float(ms.value), float(sm.value)


# ---- scene files in the reference's own formats ------------------------------------------------------------------------------

def write_obj(path, tris):
    """Triangle soup [n,9] -> ASCII .obj the reference's tinyobj loader reads (one v-triple + f per triangle)."""
    t = np.asarray(tris, np.float32).reshape(-1, 3)
    with open(path, "w") as f:
        # %.9g round-trips float32 exactly
        f.write("".join("v %.9g %.9g %.9g\n" % (a, b, c) for a, b, c in t))
        n = t.shape[0] // 3
        f.write("".join("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3) for i in range(n)))


def write_scene(path, entries):
    """entries: list of (type, path, center). The file name must contain 'geometry' for the reference to pick Geometry mode
    (mode_from_scene, src/common_host.cu:146-166)."""
    assert "geometry" in os.path.basename(path)
    with open(path, "w") as f:
        json.dump({"geometry": [{"center": [float(x) for x in c], "path": p, "type": t} for t, p, c in entries]}, f)


def write_snapshot_py(path, config_json_path, params_fp16, density_grid, aabb_scale=1, camera_cols=None, rel_focal=None):
    """A `.ingp` snapshot produced WITHOUT the product library (python msgpack + gzip, the container zstr writes): network config
    JSON + fp16 parameter blob + fp16 density grid, the members Testbed::load_snapshot requires (src/testbed.cu:5285-5463).
    Used by bench.py's reference arm so that nothing of the product touches the reference's inputs."""
    import gzip
    import msgpack
    cfg = json.load(open(config_json_path))
    cam = np.asarray(camera_cols if camera_cols is not None else [1, 0, 0, 0, -1, 0, 0, 0, -1, 0.5, 0.5, 2.0], np.float32).reshape(4, 3)
    rf = [1.0572856, 1.0572856] if rel_focal is None else [float(rel_focal[0]), float(rel_focal[1])]
    lo, hi = 0.5 - 0.5 * aabb_scale, 0.5 + 0.5 * aabb_scale
    cfg["snapshot"] = {
        "version": 1, "mode": "nerf", "n_params": int(params_fp16.size), "params_type": "__half",
        "params_binary": np.ascontiguousarray(params_fp16, np.float16).tobytes(),
        "density_grid_size": 128, "density_grid_binary": np.ascontiguousarray(density_grid, np.float32).astype(np.float16).tobytes(),
        "nerf": {"aabb_scale": int(aabb_scale), "rgb": {"rays_per_batch": 4096, "measured_batch_size": 0, "measured_batch_size_before_compaction": 0}},
        "training_step": 0, "loss": 0.0, "aabb": {"min": [lo] * 3, "max": [hi] * 3}, "render_aabb": {"min": [lo] * 3, "max": [hi] * 3},
        "up_dir": [0.0, 1.0, 0.0], "sun_dir": [0.57735026919] * 3, "exposure": 0.0, "background_color": [0.0, 0.0, 0.0, 1.0],
        "camera": {"matrix": [[float(cam[c][r]) for c in range(4)] for r in range(3)], "fov_axis": 1, "relative_focal_length": rf,
                   "screen_center": [0.5, 0.5], "zoom": 1.0},
    }
    with open(path, "wb") as f:
        f.write(gzip.compress(msgpack.packb(cfg, use_bin_type=True), compresslevel=0))


def write_png_rgba8(path, rgba8):
    """Minimal PNG writer (RGBA8, non-interlaced, filter type 0) - enough for stb_image / the product's own PNG reader."""
    import struct
    import zlib
    h, w, _ = rgba8.shape
    raw = np.concatenate([np.zeros((h, 1), np.uint8), np.ascontiguousarray(rgba8, np.uint8).reshape(h, w * 4)], 1).tobytes()

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def write_nerf_dataset(directory, images, cams, focal_px, aabb_scale=1):
    """A NeRF dataset in the reference's format (transforms.json + 8-bit sRGB PNG frames with straight alpha): images float32
    [n][h][w][4] linear RGB + alpha, cams [n][12] NGP-space columns {right, down, forward, position}. The transform matrices are the
    inverse of nerf_matrix_to_ngp (scale 0.33, offset 0.5, axis cycle), so the loader lands on exactly these cameras."""
    os.makedirs(directory, exist_ok=True)
    n, h, w, _ = images.shape
    rgb = images[..., :3]
    srgb = np.where(rgb < 0.0031308, 12.92 * rgb, 1.055 * np.power(np.maximum(rgb, 1e-12), 1 / 2.4) - 0.055)
    rgba8 = np.concatenate([np.round(np.clip(srgb, 0, 1) * 255), np.round(images[..., 3:4] * 255)], -1).astype(np.uint8)
    frames = []
    for i in range(n):
        write_png_rgba8(os.path.join(directory, f"r_{i}.png"), rgba8[i])
        c = np.asarray(cams[i], np.float64).reshape(4, 3)
        col = np.stack([c[:, 2], c[:, 0], c[:, 1]], 1)   # undo the axis cycle
        col[1] *= -1; col[2] *= -1
        col[3] = (col[3] - 0.5) / 0.33
        m = np.eye(4); m[:3, :4] = col.T
        frames.append({"file_path": f"r_{i}", "transform_matrix": m.tolist()})
    meta = {"camera_angle_x": float(2 * np.arctan(0.5 * w / focal_px)), "aabb_scale": int(aabb_scale), "frames": frames}
    with open(os.path.join(directory, "transforms.json"), "w") as f:
        json.dump(meta, f)
    return os.path.join(directory, "transforms.json")
