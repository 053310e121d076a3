"""ctypes binding of libhn_b200.so (include/hybrid_nerf.h) shaped like the reference's ``pyngp.Testbed``
(src/python_api.cu:263-733): same method names and argument meaning for the render path, errors raised as
RuntimeError (pybind11 translates the reference's std::runtime_error the same way).

There is no CPU path behind this module: importing it without the built CUDA library raises ImportError, and every
computing call fails without a B200.
"""
from __future__ import annotations

import ctypes as C
import enum
import math
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhn_b200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(make -C <package>/csrc). This package has no CPU fallback."
    )

_lib = C.CDLL(LIB_PATH)

_f32p = C.POINTER(C.c_float)
_u8p = C.POINTER(C.c_uint8)
_u16p = C.POINTER(C.c_uint16)
_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)


class NetworkInfo(C.Structure):
    _fields_ = [
        ("n_levels", C.c_uint32), ("n_features_per_level", C.c_uint32), ("log2_hashmap_size", C.c_uint32), ("base_resolution", C.c_uint32),
        ("per_level_scale", C.c_float), ("density_hidden_layers", C.c_uint32), ("rgb_hidden_layers", C.c_uint32),
        ("n_params", C.c_uint64), ("grid_param_offset", C.c_uint64),
        ("level_offsets", C.c_uint32 * 17), ("level_scales", C.c_float * 16), ("level_resolutions", C.c_uint32 * 16),
    ]


class FrameStats(C.Structure):
    _fields_ = [
        ("n_samples", C.c_uint64), ("n_samples_composited", C.c_uint64), ("n_rays_alive", C.c_uint32), ("n_passes", C.c_uint32),
        ("n_kernel_launches", C.c_uint32), ("ms_mesh", C.c_float), ("ms_raygen", C.c_float), ("ms_march", C.c_float),
        ("ms_eval", C.c_float), ("ms_composite", C.c_float), ("ms_total", C.c_float),
        ("ms_envlight", C.c_float), ("n_env_hits", C.c_uint32), ("n_light_rays", C.c_uint64), ("n_light_rays_volume", C.c_uint64),
        ("n_light_samples", C.c_uint64), ("n_bvh_stack_overflows", C.c_uint32), ("pad_", C.c_uint32),
    ]


class TrainStats(C.Structure):
    _fields_ = [
        ("step", C.c_uint32), ("rays_per_batch", C.c_uint32), ("n_rays_last", C.c_uint32), ("n_samples_last", C.c_uint32),
        ("loss_last", C.c_float), ("learning_rate", C.c_float), ("ms_sample", C.c_float), ("ms_forward", C.c_float), ("ms_loss", C.c_float),
        ("ms_backward", C.c_float), ("ms_optimizer", C.c_float),
    ]


# every symbol include/hybrid_nerf.h declares (tests check that the library exports all of them)
NOALLOC_MAX_POSITIONS = 65536  # HN_NOALLOC_MAX_POSITIONS (include/hybrid_nerf.h): positions per hn_op_encode_noalloc call

EXPORTS = [
    "hn_create", "hn_destroy", "hn_last_error", "hn_version", "hn_load_network_config", "hn_load_network_config_string",
    "hn_set_aabb_scale", "hn_get_network_info", "hn_set_params", "hn_load_snapshot", "hn_save_snapshot", "hn_set_density_grid", "hn_set_occupancy_bitfield",
    "hn_get_occupancy_bitfield", "hn_load_scene", "hn_add_mesh", "hn_load_mesh_file", "hn_build_scene", "hn_clear_scene",
    "hn_get_scene_aabb", "hn_set_option", "hn_get_option", "hn_render_frame", "hn_render_to_host", "hn_set_camera",
    "hn_set_tile_shard", "hn_tile_count", "hn_pack_tiles", "hn_unpack_tiles", "hn_get_frame_stats", "hn_op_init_rays", "hn_op_mesh_distance",
    "hn_op_march_all", "hn_op_encode", "hn_op_network", "hn_op_trace", "hn_set_envmap", "hn_op_read_envmap", "hn_op_env_light_dirs",
    "hn_op_transmittance", "hn_train_set_dataset", "hn_train_init", "hn_train_accumulate", "hn_train_gradient_buffer", "hn_train_apply",
    "hn_train_step", "hn_get_train_stats", "hn_op_train_gradients", "hn_update_density_grid", "hn_mark_untrained_density_grid",
    "hn_get_density_grid", "hn_load_training_data", "hn_shared_alloc", "hn_shared_open", "hn_shared_release", "hn_scatter_tiles", "hn_host_register", "hn_host_unregister",
    "hn_render_probe", "hn_set_probe", "hn_op_probe_rays", "hn_set_stream", "hn_op_encode_noalloc",
    "hn_get_error_map", "hn_op_error_map_cdf", "hn_op_sample_training_pixels", "hn_op_error_map_deposit",
    "hn_set_hidden_area_mask", "hn_set_distortion_map", "hn_unwarp_foveated", "hn_op_foveation", "hn_train_set_image_metadata", "hn_decode_image_file", "hn_train_set_depth", "hn_op_train_gradients_depth",
]

_lib.hn_last_error.restype = C.c_char_p
_lib.hn_version.restype = C.c_char_p
_lib.hn_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
_lib.hn_destroy.argtypes = [C.c_void_p]
_lib.hn_last_error.argtypes = [C.c_void_p]
_lib.hn_set_params.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
_lib.hn_set_occupancy_bitfield.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
_lib.hn_get_occupancy_bitfield.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
_lib.hn_add_mesh.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
_lib.hn_render_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int,
                                 C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_render_to_host.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
_lib.hn_pack_tiles.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_unpack_tiles.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_set_stream.argtypes = [C.c_void_p, C.c_void_p]
_lib.hn_render_probe.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
_lib.hn_set_probe.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
_lib.hn_op_probe_rays.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_shared_alloc.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.c_void_p]
_lib.hn_shared_open.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
_lib.hn_shared_release.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
_lib.hn_scatter_tiles.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_host_register.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
_lib.hn_host_unregister.argtypes = [C.c_void_p, C.c_void_p]
_lib.hn_op_init_rays.argtypes = [C.c_void_p] + [C.c_void_p] * 5 + [C.c_uint32, C.c_int, C.c_int] + [C.c_void_p] * 4
_lib.hn_op_march_all.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 6 + [C.c_uint32]
_lib.hn_op_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
_lib.hn_train_set_depth.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int]
_lib.hn_op_train_gradients_depth.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 6 + [C.c_float, C.c_int, C.c_int, C.c_float] + [C.c_void_p] * 3
_lib.hn_decode_image_file.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t]
_lib.hn_train_set_image_metadata.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_set_hidden_area_mask.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
_lib.hn_set_distortion_map.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
_lib.hn_unwarp_foveated.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_op_foveation.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_get_error_map.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
_lib.hn_op_error_map_cdf.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 5
_lib.hn_op_sample_training_pixels.argtypes = [C.c_void_p] + [C.c_uint32] * 6 + [C.c_int, C.c_int] + [C.c_void_p] * 6
_lib.hn_op_error_map_deposit.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
_lib.hn_op_encode_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
_lib.hn_op_network.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
_lib.hn_op_trace.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 5
_lib.hn_op_mesh_distance.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
for _n in ("hn_load_network_config", "hn_load_network_config_string", "hn_load_scene", "hn_load_snapshot", "hn_save_snapshot"):
    getattr(_lib, _n).argtypes = [C.c_void_p, C.c_char_p]
_lib.hn_load_mesh_file.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
_lib.hn_set_aabb_scale.argtypes = [C.c_void_p, C.c_int]
_lib.hn_get_network_info.argtypes = [C.c_void_p, C.POINTER(NetworkInfo)]
_lib.hn_set_density_grid.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
_lib.hn_build_scene.argtypes = [C.c_void_p]
_lib.hn_clear_scene.argtypes = [C.c_void_p]
_lib.hn_get_scene_aabb.argtypes = [C.c_void_p, C.c_void_p]
_lib.hn_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int]
_lib.hn_get_option.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int]
_lib.hn_set_camera.argtypes = [C.c_void_p, C.c_void_p]
_lib.hn_set_tile_shard.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
_lib.hn_tile_count.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, _u32p]
_lib.hn_get_frame_stats.argtypes = [C.c_void_p, C.POINTER(FrameStats)]
_lib.hn_train_set_dataset.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
_lib.hn_train_init.argtypes = [C.c_void_p, C.c_uint32]
_lib.hn_train_accumulate.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_float)]
_lib.hn_train_gradient_buffer.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
_lib.hn_train_apply.argtypes = [C.c_void_p, C.c_float]
_lib.hn_train_step.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_float)]
_lib.hn_get_train_stats.argtypes = [C.c_void_p, C.POINTER(TrainStats)]
_lib.hn_op_train_gradients.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 5 + [C.c_int, C.c_float, C.POINTER(C.c_float), C.c_void_p, C.c_void_p]
_lib.hn_load_training_data.argtypes = [C.c_void_p, C.c_char_p]
_lib.hn_update_density_grid.argtypes = [C.c_void_p, C.c_float, C.c_uint32, C.c_uint32]
_lib.hn_mark_untrained_density_grid.argtypes = [C.c_void_p]
_lib.hn_get_density_grid.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
_lib.hn_set_envmap.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
_lib.hn_op_read_envmap.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
_lib.hn_op_env_light_dirs.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
_lib.hn_op_transmittance.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]


def cudart_memcpy_d2h(out: np.ndarray, dev_ptr: int) -> int:
    """Device -> host copy of out.nbytes bytes through the CUDA runtime the library is linked against (tests, tooling)."""
    try:
        rt = C.CDLL("libcudart.so.12")
    except OSError:
        rt = C.CDLL("/usr/local/cuda/lib64/libcudart.so.12")
    rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    return int(rt.cudaMemcpy(out.ctypes.data_as(C.c_void_p), C.c_void_p(dev_ptr), out.nbytes, 2))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def camera_to_cols(m):
    """3x4 row-major camera-to-world matrix (pyngp's set_nerf_camera_matrix argument) -> the ABI's 4 columns x vec3."""
    m = np.asarray(m, dtype=np.float32).reshape(3, 4)
    return np.ascontiguousarray(m.T).reshape(12)


def decode_image_file(path: str) -> np.ndarray:
    """the dataset loader's image decoder (PNG / baseline JPEG -> uint8 [h][w][4]); needs no GPU"""
    w, h = C.c_int(0), C.c_int(0)
    err = C.create_string_buffer(512)
    if _lib.hn_decode_image_file(os.fsencode(path), C.byref(w), C.byref(h), None, 0, err, 512):
        raise RuntimeError(err.value.decode())
    out = np.empty((h.value, w.value, 4), np.uint8)
    if _lib.hn_decode_image_file(os.fsencode(path), C.byref(w), C.byref(h), _ptr(out), out.size, err, 512):
        raise RuntimeError(err.value.decode())
    return out


class RenderMode(enum.IntEnum):
    AO = 0
    Shade = 1
    ShadeNerf = 2
    ShadeEnvMap = 3
    ShadeGridEnvMap = 4
    Normals = 5
    Positions = 6
    Depth = 7
    Distortion = 8
    Cost = 9
    Slice = 10
    EncodingVis = 12


class Testbed:
    """Render-path subset of ``pyngp.Testbed``."""

    def __init__(self, mode: str = "nerf", device: int = 0):
        self._h = C.c_void_p()
        rc = _lib.hn_create(C.byref(self._h), int(device))
        if rc != 0:
            msg = _lib.hn_last_error(self._h).decode() if self._h else "hn_create failed"
            if self._h:
                _lib.hn_destroy(self._h)
                self._h = C.c_void_p()
            raise RuntimeError(msg)
        self.mode = mode

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.hn_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(_lib.hn_last_error(self._h).decode())

    # ---- network / scene (python_api.cu:429-441)
    def reload_network_from_file(self, path: str):
        self._check(_lib.hn_load_network_config(self._h, os.fsencode(path)))

    def reload_network_from_json(self, text: str):
        self._check(_lib.hn_load_network_config_string(self._h, text.encode()))

    def load_snapshot(self, path: str):
        """pyngp.Testbed.load_snapshot (python_api.cu:436): .ingp / .msgpack written by the reference or by save_snapshot."""
        self._check(_lib.hn_load_snapshot(self._h, os.fsencode(path)))

    def save_snapshot(self, path: str, include_optimizer_state: bool = False):
        """pyngp.Testbed.save_snapshot(path, include_optimizer_state) (python_api.cu:437): with the flag the optimizer stack
        (Adam moments, per-parameter step counters, EMA weights) is serialised as tcnn's Trainer::serialize(true) does."""
        self.set_option("snapshot_optimizer_state", float(bool(include_optimizer_state)))
        self._check(_lib.hn_save_snapshot(self._h, os.fsencode(path)))

    def set_aabb_scale(self, s: int):
        self._check(_lib.hn_set_aabb_scale(self._h, int(s)))

    def network_info(self) -> NetworkInfo:
        info = NetworkInfo()
        self._check(_lib.hn_get_network_info(self._h, C.byref(info)))
        return info

    def set_params(self, blob: np.ndarray):
        blob = np.ascontiguousarray(blob)
        if blob.dtype == np.float16:
            dt = 0
        elif blob.dtype == np.float32:
            dt = 1
        else:
            raise TypeError("params must be float16 or float32")
        self._check(_lib.hn_set_params(self._h, _ptr(blob), blob.size, dt))

    def set_density_grid(self, grid: np.ndarray, n_cascades: int = 1):
        g = _f32(grid)
        assert g.size == 128 ** 3 * n_cascades
        self._check(_lib.hn_set_density_grid(self._h, _ptr(g), n_cascades))

    def set_occupancy_bitfield(self, bits: np.ndarray):
        b = np.ascontiguousarray(bits, dtype=np.uint8)
        self._check(_lib.hn_set_occupancy_bitfield(self._h, _ptr(b), b.size))

    def get_occupancy_bitfield(self) -> np.ndarray:
        out = np.zeros(128 ** 3 // 8 * 8, dtype=np.uint8)
        self._check(_lib.hn_get_occupancy_bitfield(self._h, _ptr(out), out.size))
        return out

    def load_file(self, path: str):
        """Scene json (Geometry mode) or mesh file, like Testbed::load_file (src/testbed.cu:319-389)."""
        if path.endswith(".json"):
            self._check(_lib.hn_load_scene(self._h, os.fsencode(path)))
        elif path.endswith((".ingp", ".msgpack")):
            self.load_snapshot(path)
        else:
            c = _f32([0, 0, 0])
            self._check(_lib.hn_load_mesh_file(self._h, os.fsencode(path), _ptr(c)))
            self.build_scene()

    load_training_data = load_file

    def add_mesh(self, triangles: np.ndarray, center=(0.0, 0.0, 0.0), normalize: bool = False):
        t = _f32(triangles).reshape(-1, 9)
        c = _f32(center)
        self._check(_lib.hn_add_mesh(self._h, _ptr(t), t.shape[0], _ptr(c), int(normalize)))

    def build_scene(self):
        self._check(_lib.hn_build_scene(self._h))

    def clear_scene(self):
        self._check(_lib.hn_clear_scene(self._h))

    def scene_aabb(self) -> np.ndarray:
        out = np.zeros(6, dtype=np.float32)
        self._check(_lib.hn_get_scene_aabb(self._h, _ptr(out)))
        return out

    # ---- options
    def set_option(self, name: str, value):
        v = _f32(np.atleast_1d(value))
        self._check(_lib.hn_set_option(self._h, name.encode(), _ptr(v), v.size))

    def get_option(self, name: str, n: int = 1):
        v = np.zeros(n, dtype=np.float32)
        self._check(_lib.hn_get_option(self._h, name.encode(), _ptr(v), n))
        return v if n > 1 else float(v[0])

    @property
    def fov(self) -> float:  # python_api.cu: fov in degrees on fov_axis
        f = self.get_option("relative_focal_length", 2)[int(self.get_option("fov_axis"))]
        return 2.0 * 180.0 / math.pi * math.atan(1.0 / (2.0 * f))

    @fov.setter
    def fov(self, deg: float):
        f = 0.5 / math.tan(0.5 * deg * math.pi / 180.0)
        self.set_option("relative_focal_length", [f, f])

    snap_to_pixel_centers = property(lambda s: bool(s.get_option("snap_to_pixel_centers")), lambda s, v: s.set_option("snap_to_pixel_centers", float(v)))
    exposure = property(lambda s: s.get_option("exposure"), lambda s, v: s.set_option("exposure", v))
    sun_dir = property(lambda s: s.get_option("sun_dir", 3), lambda s, v: s.set_option("sun_dir", v))
    up_dir = property(lambda s: s.get_option("up_dir", 3), lambda s, v: s.set_option("up_dir", v))
    # ERenderMode (include/neural-graphics-primitives/common.h:58-72, pyngp RenderMode); the NeRF pass builds AO, Shade*, Positions,
    # Depth, Cost and EncodingVis, the others raise at render time
    render_mode = property(lambda s: RenderMode(int(s.get_option("render_mode"))), lambda s, v: s.set_option("render_mode", float(int(v))))
    aperture_size = property(lambda s: s.get_option("aperture_size"), lambda s, v: s.set_option("aperture_size", v))
    parallax_shift = property(lambda s: s.get_option("parallax_shift", 3), lambda s, v: s.set_option("parallax_shift", v))
    background_color = property(lambda s: s.get_option("background_color", 4), lambda s, v: s.set_option("background_color", v))

    def set_nerf_camera_matrix(self, m):
        c = camera_to_cols(m)
        self._check(_lib.hn_set_camera(self._h, _ptr(c)))

    def set_camera_cols(self, cols12):
        c = _f32(cols12).reshape(12)
        self._check(_lib.hn_set_camera(self._h, _ptr(c)))

    def set_tile_shard(self, rank: int, world: int, tile_w: int = 64, tile_h: int = 32):
        self._check(_lib.hn_set_tile_shard(self._h, rank, world, tile_w, tile_h))

    # ---- rendering
    def render(self, width: int, height: int, spp: int = 1, linear: bool = True, return_depth: bool = False, out: np.ndarray | None = None):
        """pyngp.Testbed.render(width, height, spp, linear) -> float32 [h, w, 4] (python_api.cu:417-426).
        `out` may be a caller-owned (e.g. pinned) float32 [h, w, 4] array."""
        if out is None:
            out = np.empty((height, width, 4), dtype=np.float32)
        assert out.dtype == np.float32 and out.size == height * width * 4 and out.flags["C_CONTIGUOUS"]
        depth = np.empty((height, width), dtype=np.float32) if return_depth else None
        self._check(_lib.hn_render_to_host(self._h, width, height, spp, int(linear), _ptr(out), _ptr(depth) if return_depth else None))
        return (out, depth) if return_depth else out

    def render_frame(self, cam0, cam1, width, height, frame_ptr: int, depth_ptr: int, spp_index: int = 0, stream: int = 0,
                     screen_center=(0.5, 0.5), relative_focal_length=None, rolling_shutter=(0.0, 0.0, 0.0, 1.0)):
        """Testbed::render_frame into caller-owned DEVICE buffers (raw pointers, e.g. torch.Tensor.data_ptr())."""
        c0, c1 = _f32(cam0).reshape(12), _f32(cam1).reshape(12)
        sc = _f32(screen_center)
        rf = _f32(relative_focal_length if relative_focal_length is not None else self.get_option("relative_focal_length", 2))
        rs = _f32(rolling_shutter)
        self._check(_lib.hn_render_frame(self._h, _ptr(c0), _ptr(c1), _ptr(sc), _ptr(rf), _ptr(rs), spp_index, width, height,
                                         C.c_void_p(frame_ptr), C.c_void_p(depth_ptr), C.c_void_p(stream)))

    def frame_stats(self) -> FrameStats:
        s = FrameStats()
        self._check(_lib.hn_get_frame_stats(self._h, C.byref(s)))
        return s

    def tile_count(self, width, height, rank, world) -> int:
        n = C.c_uint32()
        self._check(_lib.hn_tile_count(self._h, width, height, rank, world, C.byref(n)))
        return n.value

    def pack_tiles(self, width, height, rank, world, frame_ptr, depth_ptr, packed_ptr, stream=0):
        self._check(_lib.hn_pack_tiles(self._h, width, height, rank, world, C.c_void_p(frame_ptr), C.c_void_p(depth_ptr), C.c_void_p(packed_ptr), C.c_void_p(stream)))

    def unpack_tiles(self, width, height, rank, world, packed_ptr, frame_ptr, depth_ptr, stream=0):
        self._check(_lib.hn_unpack_tiles(self._h, width, height, rank, world, C.c_void_p(packed_ptr), C.c_void_p(frame_ptr), C.c_void_p(depth_ptr), C.c_void_p(stream)))

    # ---- frame gather over peer memory (one process per GPU)
    def shared_alloc(self, n_bytes: int):
        """-> (device pointer, 64-byte CUDA IPC handle) of a buffer other ranks of this node can map with shared_open"""
        ptr = C.c_void_p(0)
        handle = (C.c_ubyte * 64)()
        self._check(_lib.hn_shared_alloc(self._h, n_bytes, C.byref(ptr), handle))
        return int(ptr.value), bytes(handle)

    def shared_open(self, handle: bytes) -> int:
        ptr = C.c_void_p(0)
        buf = (C.c_ubyte * 64).from_buffer_copy(handle)
        self._check(_lib.hn_shared_open(self._h, buf, C.byref(ptr)))
        return int(ptr.value)

    def shared_release(self, ptr: int, owner: bool):
        self._check(_lib.hn_shared_release(self._h, C.c_void_p(ptr), int(owner)))

    def scatter_tiles(self, width, height, rank, world, frame_ptr, depth_ptr, frame_remote, depth_remote, stream=0):
        """copy this rank's tiles of the local frame / depth into the root rank's buffers (mapped with shared_open) or into host memory
        (mapped with host_register; depth_remote may be 0 there: colour only)"""
        self._check(_lib.hn_scatter_tiles(self._h, width, height, rank, world, C.c_void_p(frame_ptr), C.c_void_p(depth_ptr),
                                          C.c_void_p(frame_remote), C.c_void_p(depth_remote or None), C.c_void_p(stream)))

    def host_register(self, host_ptr: int, n_bytes: int) -> int:
        """page-lock + map host memory (e.g. tiles.SharedHostFrame) -> pointer scatter_tiles accepts as frame_remote / depth_remote"""
        ptr = C.c_void_p(0)
        self._check(_lib.hn_host_register(self._h, C.c_void_p(host_ptr), n_bytes, C.byref(ptr)))
        return int(ptr.value)

    def host_unregister(self, host_ptr: int):
        self._check(_lib.hn_host_unregister(self._h, C.c_void_p(host_ptr)))

    # ---- operator-level entry points (parity tests)
    def op_init_rays(self, cam0, cam1, width, height, spp_index=0, screen_center=(0.5, 0.5), relative_focal_length=None, rolling_shutter=(0, 0, 0, 1)):
        n = width * height
        o = np.empty((n, 3), np.float32); d = np.empty((n, 3), np.float32); t = np.empty(n, np.float32); alive = np.empty(n, np.uint8)
        c0, c1 = _f32(cam0).reshape(12), _f32(cam1).reshape(12)
        sc = _f32(screen_center)
        rf = _f32(relative_focal_length if relative_focal_length is not None else self.get_option("relative_focal_length", 2))
        rs = _f32(rolling_shutter)
        self._check(_lib.hn_op_init_rays(self._h, _ptr(c0), _ptr(c1), _ptr(sc), _ptr(rf), _ptr(rs), spp_index, width, height, _ptr(o), _ptr(d), _ptr(t), _ptr(alive)))
        return o, d, t, alive

    # ---- VR pieces of the camera model (Testbed::m_views[i].foveation, render_buffer.hidden_area_mask, m_distortion)
    def set_hidden_area_mask(self, mask):
        """uint8 [h][w], 0 = hidden (no ray); None removes it"""
        if mask is None:
            self._check(_lib.hn_set_hidden_area_mask(self._h, None, 0, 0)); return
        m = np.ascontiguousarray(mask, np.uint8)
        self._check(_lib.hn_set_hidden_area_mask(self._h, _ptr(m), m.shape[1], m.shape[0]))

    def set_distortion_map(self, grid):
        """float32 [h][w][2] offsets added to the camera-space direction of NeRF rays (bilinear); None removes it"""
        if grid is None:
            self._check(_lib.hn_set_distortion_map(self._h, None, 0, 0)); return
        g = _f32(grid)
        assert g.ndim == 3 and g.shape[2] == 2
        self._check(_lib.hn_set_distortion_map(self._h, _ptr(g), g.shape[1], g.shape[0]))

    def op_foveation(self, foveation):
        f = _f32(foveation).reshape(6); out = np.empty((2, 12), np.float32)
        self._check(_lib.hn_op_foveation(self._h, _ptr(f), _ptr(out)))
        return out

    def unwarp_foveated(self, frame, foveation, out_res=None, depth=None):
        """frame float32 [h][w][4] rendered with option `foveation` -> the un-warped image at out_res (w, h); host arrays in / out"""
        import torch
        f = _f32(foveation).reshape(6)
        fin = torch.from_numpy(_f32(frame)).cuda()
        h, w = fin.shape[0], fin.shape[1]
        ow, oh = (w, h) if out_res is None else out_res
        fout = torch.empty(oh, ow, 4, device="cuda")
        din = torch.from_numpy(_f32(depth)).cuda() if depth is not None else None
        dout = torch.empty(oh, ow, device="cuda") if depth is not None else None
        torch.cuda.synchronize()
        self._check(_lib.hn_unwarp_foveated(self._h, _ptr(f), w, h, fin.data_ptr(), din.data_ptr() if din is not None else None, ow, oh, fout.data_ptr(),
                                            dout.data_ptr() if dout is not None else None, None))
        torch.cuda.synchronize()
        return (fout.cpu().numpy(), dout.cpu().numpy()) if depth is not None else fout.cpu().numpy()

    def op_march_all(self, origins, dirs, t0, alive, max_per_ray=0):
        o, d, t = _f32(origins), _f32(dirs), _f32(t0)
        a = np.ascontiguousarray(alive, dtype=np.uint8)
        n = t.size
        counts = np.empty(n, np.uint32)
        samples = np.empty((n, max_per_ray, 4), np.float32) if max_per_ray else None
        self._check(_lib.hn_op_march_all(self._h, n, _ptr(o), _ptr(d), _ptr(t), _ptr(a), _ptr(counts), _ptr(samples) if max_per_ray else None, max_per_ray))
        return counts, samples

    def op_encode(self, positions):
        p = _f32(positions).reshape(-1, 3)
        out = np.empty((p.shape[0], 32), np.float16)
        self._check(_lib.hn_op_encode(self._h, _ptr(p), p.shape[0], _ptr(out)))
        return out

    def op_encode_noalloc(self, positions):
        """hn_op_encode through the library's static device scratch: no device allocation, at most NOALLOC_MAX_POSITIONS per call"""
        p = _f32(positions).reshape(-1, 3)
        out = np.empty((p.shape[0], 32), np.float16)
        self._check(_lib.hn_op_encode_noalloc(self._h, _ptr(p), p.shape[0], _ptr(out)))
        return out

    def op_network(self, positions_dt, dirs):
        p = _f32(positions_dt).reshape(-1, 4)
        d = _f32(dirs).reshape(-1, 3)
        assert p.shape[0] == d.shape[0] * 32
        out = np.empty((p.shape[0], 4), np.float16)
        self._check(_lib.hn_op_network(self._h, _ptr(p), _ptr(d), p.shape[0], _ptr(out)))
        return out

    # ---- a-21 training (pyngp: Testbed.train / .frame with shall_train, .loss, .training_step; src/python_api.cu:427-440,584-590)
    def set_training_data(self, images: np.ndarray, cameras: np.ndarray, focal_px, principal=(0.5, 0.5)):
        """images float32 [n][h][w][4] (linear RGB, straight alpha), cameras [n][12] (mat4x3 columns), pinhole intrinsics."""
        im = _f32(images)
        cams = _f32(cameras).reshape(-1, 12)
        if im.ndim != 4 or im.shape[3] != 4 or im.shape[0] != cams.shape[0]:
            raise RuntimeError("images must be [n][h][w][4] with one camera each")
        f = _f32(np.broadcast_to(np.asarray(focal_px, np.float32), (2,)).copy())
        pp = _f32(principal)
        self._check(_lib.hn_train_set_dataset(self._h, _ptr(im), im.shape[0], im.shape[2], im.shape[1], _ptr(cams), _ptr(f), _ptr(pp)))

    def set_image_metadata(self, focal_px, principal=None, lens_mode=None, lens_params=None):
        """per-image cameras after set_training_data: focal_px [n][2], principal [n][2], lens_mode [n] (ELensMode), lens_params [n][7]"""
        f = _f32(focal_px).reshape(-1, 2)
        pp = _f32(principal).reshape(-1, 2) if principal is not None else None
        lm = np.ascontiguousarray(lens_mode, np.int32).reshape(-1) if lens_mode is not None else None
        lp = _f32(lens_params).reshape(-1, 7) if lens_params is not None else None
        self._check(_lib.hn_train_set_image_metadata(self._h, f.shape[0], _ptr(f), _ptr(pp) if pp is not None else None, _ptr(lm) if lm is not None else None,
                                                     _ptr(lp) if lp is not None else None))

    def load_training_data(self, path: str):
        """pyngp Testbed.load_training_data (src/python_api.cu:409): transforms.json (or its folder) with PNG frames."""
        self._check(_lib.hn_load_training_data(self._h, path.encode()))

    def reset_trainer(self, seed: int = 1337):
        self._check(_lib.hn_train_init(self._h, seed))

    def train(self, n_steps: int = 1) -> float:
        loss = C.c_float(0)
        self._check(_lib.hn_train_step(self._h, n_steps, C.byref(loss)))
        return float(loss.value)

    def train_accumulate(self, n_rays: int = 0) -> float:
        loss = C.c_float(0)
        self._check(_lib.hn_train_accumulate(self._h, n_rays, C.byref(loss)))
        return float(loss.value)

    def train_gradient_buffer(self):
        """(device pointer, number of floats) of the fp32 gradient buffer: all-reduce it between accumulate and apply."""
        ptr = C.c_void_p(0); n = C.c_size_t(0)
        self._check(_lib.hn_train_gradient_buffer(self._h, C.byref(ptr), C.byref(n)))
        return int(ptr.value), int(n.value)

    def set_stream(self, stream: int = 0):
        """run the training / set-up calls on the caller's CUDA stream (0 / None = the context's own)"""
        self._check(_lib.hn_set_stream(self._h, C.c_void_p(stream or 0)))

    def train_apply(self, grad_scale: float = 1.0):
        self._check(_lib.hn_train_apply(self._h, grad_scale))

    def op_mesh_distance(self, positions, mode: int = 0, max_distance: float = 10.0, return_triangles: bool = False):
        """TriangleBvh unsigned / watertight-signed distance (src/triangle_bvh.cu:285-297) for an (n, 3) array of positions"""
        p = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        d = np.empty(p.shape[0], np.float32)
        tri = np.empty(p.shape[0], np.int32) if return_triangles else None
        self._check(_lib.hn_op_mesh_distance(self._h, C.c_uint32(p.shape[0]), _ptr(p), C.c_int(mode), C.c_float(max_distance), _ptr(d),
                                             _ptr(tri) if tri is not None else None))
        return (d, tri) if return_triangles else d

    def update_density_grid(self, decay: float = 0.95, n_uniform: int = 128 ** 3, n_nonuniform: int = 0):
        self._check(_lib.hn_update_density_grid(self._h, decay, n_uniform, n_nonuniform))

    def mark_untrained_density_grid(self):
        self._check(_lib.hn_mark_untrained_density_grid(self._h))

    def get_density_grid(self, n_cascades: int = 1) -> np.ndarray:
        out = np.empty(128 ** 3 * n_cascades, np.float32)
        self._check(_lib.hn_get_density_grid(self._h, _ptr(out), out.size))
        return out

    def train_stats(self) -> TrainStats:
        st = TrainStats()
        _lib.hn_get_train_stats(self._h, C.byref(st))
        return st

    @property
    def loss(self) -> float:
        return float(self.train_stats().loss_last)

    @property
    def training_step(self) -> int:
        return int(self.train_stats().step)

    def op_train_gradients(self, origins, dirs, t_start, targets, backgrounds, loss_type=1, loss_scale=128.0):
        o, d, t = _f32(origins).reshape(-1, 3), _f32(dirs).reshape(-1, 3), _f32(t_start).reshape(-1)
        tg, bg = _f32(targets).reshape(-1, 3), _f32(backgrounds).reshape(-1, 3)
        n = o.shape[0]
        grads = np.empty(int(self.network_info().n_params), np.float32)
        steps = np.empty(n, np.uint32)
        loss = C.c_float(0)
        self._check(_lib.hn_op_train_gradients(self._h, n, _ptr(o), _ptr(d), _ptr(t), _ptr(tg), _ptr(bg), int(loss_type), float(loss_scale),
                                               C.byref(loss), _ptr(grads), _ptr(steps)))
        return float(loss.value), grads, steps

    # ---- error map (m_nerf.training.error_map; nerf.h:104-118)
    def error_map(self):
        """-> float32 [n_images][res_y][res_x]: mean training loss deposited where the rays of the current accumulation period landed"""
        res = (C.c_int * 2)()
        n = C.c_uint32(0)
        self._check(_lib.hn_get_error_map(self._h, res, C.byref(n), None, 0))
        out = np.empty((int(n.value), int(res[1]), int(res[0])), np.float32)
        self._check(_lib.hn_get_error_map(self._h, res, C.byref(n), _ptr(out), out.size))
        return out

    def op_error_map_cdf(self, data):
        d = _f32(data)
        assert d.ndim == 3
        n, h, w = d.shape
        cx, cy, ci, pi = np.empty_like(d), np.empty((n, h), np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        self._check(_lib.hn_op_error_map_cdf(self._h, n, h, w, _ptr(d), _ptr(cx), _ptr(cy), _ptr(ci), _ptr(pi)))
        return cx, cy, ci, pi

    def op_sample_training_pixels(self, n_rays, n_images, seed=1337, step=0, ray_offset=0, n_rays_total=None, cdf_x_cond_y=None, cdf_y=None, cdf_img=None):
        total = int(n_rays_total if n_rays_total is not None else n_rays)
        cx = _f32(cdf_x_cond_y) if cdf_x_cond_y is not None else None
        cy = _f32(cdf_y) if cdf_y is not None else None
        ci = _f32(cdf_img) if cdf_img is not None else None
        rx, ry = (cx.shape[2], cx.shape[1]) if cx is not None else (0, 0)
        img, uv, pdf = np.empty(n_rays, np.uint32), np.empty((n_rays, 2), np.float32), np.empty(n_rays, np.float32)
        self._check(_lib.hn_op_sample_training_pixels(self._h, int(n_rays), int(ray_offset), total, int(seed), int(step), int(n_images), rx, ry,
                                                      _ptr(cx) if cx is not None else None, _ptr(cy) if cy is not None else None, _ptr(ci) if ci is not None else None,
                                                      _ptr(img), _ptr(uv), _ptr(pdf)))
        return img, uv, pdf

    def op_error_map_deposit(self, img, uv, val, n_images, res, image_res):
        i, u, v = np.ascontiguousarray(img, np.uint32).reshape(-1), _f32(uv).reshape(-1, 2), _f32(val).reshape(-1)
        out = np.empty((int(n_images), int(res[1]), int(res[0])), np.float32)
        self._check(_lib.hn_op_error_map_deposit(self._h, i.size, _ptr(i), _ptr(u), _ptr(v), int(n_images), int(res[0]), int(res[1]), int(image_res[0]), int(image_res[1]), _ptr(out)))
        return out

    def set_training_depth(self, depth):
        """float32 [n][h][w] depth images of the training set (z-depth, <= 0 = none); None removes them. Used when option depth_supervision_lambda > 0."""
        if depth is None:
            self._check(_lib.hn_train_set_depth(self._h, None, 0, 0, 0)); return
        d = _f32(depth)
        assert d.ndim == 3
        self._check(_lib.hn_train_set_depth(self._h, _ptr(d), d.shape[0], d.shape[2], d.shape[1]))

    def op_train_gradients_depth(self, origins, dirs, t_start, targets, backgrounds, target_depth, depth_lambda, depth_loss_type=3, loss_type=1, loss_scale=128.0):
        o, d, t = _f32(origins).reshape(-1, 3), _f32(dirs).reshape(-1, 3), _f32(t_start).reshape(-1)
        tg, bg, td = _f32(targets).reshape(-1, 3), _f32(backgrounds).reshape(-1, 3), _f32(target_depth).reshape(-1)
        n = o.shape[0]
        grads = np.empty(int(self.network_info().n_params), np.float32)
        steps = np.empty(n, np.uint32)
        loss = C.c_float(0)
        self._check(_lib.hn_op_train_gradients_depth(self._h, n, _ptr(o), _ptr(d), _ptr(t), _ptr(tg), _ptr(bg), _ptr(td), float(depth_lambda), int(depth_loss_type),
                                                     int(loss_type), float(loss_scale), C.byref(loss), _ptr(grads), _ptr(steps)))
        return float(loss.value), grads, steps

    # ---- a-16 environment map (Testbed::m_envmap, testbed.h:1227-1241)
    def set_envmap(self, rgba: np.ndarray | None):
        """rgba: float32 [h][w][4] lat-long map, or None to remove it. Background of camera rays and, with option
        `env_light_samples` > 0, light source of the mesh pass."""
        if rgba is None:
            self._check(_lib.hn_set_envmap(self._h, None, 0, 0))
            return
        e = _f32(rgba)
        if e.ndim != 3 or e.shape[2] != 4:
            raise RuntimeError("envmap must have shape [h][w][4]")
        self._check(_lib.hn_set_envmap(self._h, _ptr(e), e.shape[1], e.shape[0]))

    # ---- a-16 irradiance probes (NerfTracer::init_rays_from_center* + trace_mesh; the texture and its use are defined here)
    def render_probe(self, mode: int = 0, n_theta: int = 64, n_phi: int = 128, n_origin: int = 1, origin=(0.0, 0.0, 0.0), to_host: bool = True):
        """-> float32 [n_phi][n_theta][4] premultiplied linear radiance of the NeRF on the (cos theta, phi) lattice; kept on the device
        as the light of the mesh pass (options env_light_source = 1, env_light_sampling, env_light_samples)."""
        o = _f32(origin).reshape(3)
        out = np.empty((n_phi, n_theta, 4), np.float32) if to_host else None
        self._check(_lib.hn_render_probe(self._h, int(mode), n_theta, n_phi, n_origin, _ptr(o), _ptr(out) if to_host else None))
        return out

    def set_probe(self, rgba):
        if rgba is None:
            self._check(_lib.hn_set_probe(self._h, None, 0, 0))
            return
        t = _f32(rgba)
        self._check(_lib.hn_set_probe(self._h, _ptr(t), t.shape[1], t.shape[0]))

    def op_probe_rays(self, mode, n_theta, n_phi, n_origin=1, origin=(0.0, 0.0, 0.0)):
        n = n_theta * n_phi * (n_origin * n_origin if mode == 2 else 1)
        og = _f32(origin).reshape(3)
        o = np.empty((n, 3), np.float32); d = np.empty((n, 3), np.float32); alive = np.empty(n, np.uint8)
        self._check(_lib.hn_op_probe_rays(self._h, int(mode), n_theta, n_phi, n_origin, _ptr(og), _ptr(o), _ptr(d), _ptr(alive)))
        return o, d, alive

    def op_read_envmap(self, dirs):
        d = _f32(dirs).reshape(-1, 3)
        out = np.empty((d.shape[0], 4), np.float32)
        self._check(_lib.hn_op_read_envmap(self._h, _ptr(d), d.shape[0], _ptr(out)))
        return out

    def op_env_light_dirs(self, pixel_index, spp_index, K, normal):
        n = _f32(normal).reshape(3)
        out = np.empty((K, 3), np.float32)
        self._check(_lib.hn_op_env_light_dirs(self._h, int(pixel_index), int(spp_index), int(K), _ptr(n), _ptr(out)))
        return out

    def op_transmittance(self, origins, dirs):
        o, d = _f32(origins).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
        T = np.empty(o.shape[0], np.float32); ns = np.empty(o.shape[0], np.uint32)
        self._check(_lib.hn_op_transmittance(self._h, o.shape[0], _ptr(o), _ptr(d), _ptr(T), _ptr(ns)))
        return T, ns

    def op_trace(self, origins, dirs):
        o, d = _f32(origins).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
        n = o.shape[0]
        mesh = np.empty(n, np.int32); tri = np.empty(n, np.int32); t = np.empty(n, np.float32)
        self._check(_lib.hn_op_trace(self._h, n, _ptr(o), _ptr(d), _ptr(mesh), _ptr(tri), _ptr(t)))
        return mesh, tri, t
