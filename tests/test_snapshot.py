"""Snapshots (.ingp / .msgpack, src/testbed.cu:5219-5488). CPU part: a snapshot written by INDEPENDENT implementations of
the container formats (python `msgpack` + `zlib`, real DEFLATE with dynamic Huffman blocks) must parse in the library —
without a GPU the call gets as far as the parameter upload, and the network shape it reports proves the decode. GPU part:
snapshot round trips render the identical frame."""
import ctypes as C
import gzip
import os
import zlib

import msgpack
import numpy as np
import pytest


def make_snapshot(root, synthetic, info_like, path, compress=True, params_type="__half", aabb_scale=1, level=6, container="zlib"):
    import json
    cfg = json.load(open(os.path.join(root, "configs", "nerf", "base.json")))
    blob = synthetic.random_params(info_like, 21, 1.0)
    grid = synthetic.sphere_density_grid(0.3).astype(np.float16)
    params = blob.tobytes() if params_type == "__half" else blob.astype(np.float32).tobytes()
    cfg["snapshot"] = {
        "version": 1, "mode": "nerf", "n_params": int(blob.size), "params_type": params_type, "params_binary": params,
        "density_grid_size": 128, "density_grid_binary": grid.tobytes(),
        "nerf": {"aabb_scale": aabb_scale, "rgb": {"rays_per_batch": 4096, "measured_batch_size": 0, "measured_batch_size_before_compaction": 0}},
        "training_step": 1234, "loss": 0.0123, "aabb": {"min": [0.0, 0.0, 0.0], "max": [1.0, 1.0, 1.0]},
        "render_aabb": {"min": [0.0, 0.0, 0.0], "max": [1.0, 1.0, 1.0]}, "up_dir": [0.0, 1.0, 0.0], "sun_dir": [0.0, 0.6, 0.8],
        "exposure": 0.5, "background_color": [0.0, 0.0, 0.0, 1.0],
        "camera": {"matrix": [[1.0, 0.0, 0.0, 0.5], [0.0, -1.0, 0.0, 0.5], [0.0, 0.0, -1.0, 2.25]], "fov_axis": 1,
                   "relative_focal_length": [1.2, 1.2], "screen_center": [0.5, 0.5], "zoom": 1.0},
    }
    raw = msgpack.packb(cfg, use_bin_type=True)
    with open(path, "wb") as f:
        if not compress:
            f.write(raw)
        elif container == "gzip":  # what the reference writes: zstr::ostream -> deflateInit2(.., 15 + 16, ..) = gzip member
            f.write(gzip.compress(raw, compresslevel=level))
        else:
            f.write(zlib.compress(raw, level))
    return blob, grid.astype(np.float32)


class BaseInfo:
    n_levels, n_features_per_level, density_hidden_layers, rgb_hidden_layers = 8, 4, 1, 2
    n_params = 2048 + 1024 + 2048 + 4096 + 1024 + 2920448 * 4
    grid_param_offset = 10240


def _no_gpu():
    import torch
    return not torch.cuda.is_available()


@pytest.mark.parametrize("container,level", [("gzip", 6), ("gzip", 0), ("zlib", 0), ("plain", 0)])
def test_ingp_containers_decode_on_host(hn, root, tmp_path, container, level):
    """.ingp as the reference writes it (gzip member, dynamic Huffman or level 0 = stored blocks: save_snapshot(compress=false)),
    a zlib stream, and an uncompressed .ingp (zstr::istream passes unknown magic through) all decode."""
    from importlib import import_module
    tbm = import_module(hn.__name__ + ".testbed")
    lib = tbm._lib
    h = C.c_void_p()
    lib.hn_create(C.byref(h), 0)
    path = str(tmp_path / "snap.ingp")
    make_snapshot(root, hn.synthetic, BaseInfo, path, compress=container != "plain", level=level, container=container)
    head = open(path, "rb").read(2)
    assert head == (b"\x1f\x8b" if container == "gzip" else head)
    rc = lib.hn_load_snapshot(h, os.fsencode(path))
    info = tbm.NetworkInfo()
    assert lib.hn_get_network_info(h, C.byref(info)) == 0, lib.hn_last_error(h)
    assert (info.n_levels, info.n_features_per_level, info.log2_hashmap_size, info.n_params) == (8, 4, 19, BaseInfo.n_params)
    if not _no_gpu():
        assert rc == 0, lib.hn_last_error(h)
    # a corrupted gzip trailer (CRC32) is rejected
    if container == "gzip":
        data = bytearray(open(path, "rb").read())
        data[-6] ^= 0xFF
        bad = tmp_path / "badcrc.ingp"
        bad.write_bytes(bytes(data))
        assert lib.hn_load_snapshot(h, os.fsencode(str(bad))) != 0 and b"CRC32" in lib.hn_last_error(h)
    lib.hn_destroy(h)


@pytest.mark.parametrize("ext,compress,ptype", [(".ingp", True, "__half"), (".msgpack", False, "float")])
def test_snapshot_decodes_on_host(hn, root, tmp_path, ext, compress, ptype):
    from importlib import import_module
    tbm = import_module(hn.__name__ + ".testbed")
    lib = tbm._lib
    h = C.c_void_p()
    lib.hn_create(C.byref(h), 0)
    path = str(tmp_path / ("snap" + ext))
    make_snapshot(root, hn.synthetic, BaseInfo, path, compress, ptype)
    rc = lib.hn_load_snapshot(h, os.fsencode(path))
    info = tbm.NetworkInfo()
    assert lib.hn_get_network_info(h, C.byref(info)) == 0, lib.hn_last_error(h)
    assert (info.n_levels, info.n_features_per_level, info.log2_hashmap_size, info.n_params) == (8, 4, 19, BaseInfo.n_params)
    if _no_gpu():
        assert rc != 0 and b"cuda" in lib.hn_last_error(h).lower()   # decode done, upload needs the device: no fallback
        sun = (C.c_float * 3)()
        lib.hn_get_option(h, b"sun_dir", sun, 3)
        assert [round(v, 6) for v in sun] == [0.0, 0.6, 0.8]           # snapshot members were applied before the upload
    else:
        assert rc == 0, lib.hn_last_error(h)
    # corrupt / foreign files fail with a message
    bad = tmp_path / "bad.ingp"
    bad.write_bytes(b"not a snapshot at all")   # passes through the container detection as plain bytes, fails in the MessagePack reader
    assert lib.hn_load_snapshot(h, os.fsencode(str(bad))) != 0 and len(lib.hn_last_error(h)) > 0
    trunc = tmp_path / "trunc.ingp"
    trunc.write_bytes(open(path, "rb").read()[:1000] if compress else zlib.compress(b"\x81\xa1a\x01"))
    assert lib.hn_load_snapshot(h, os.fsencode(str(trunc))) != 0
    lib.hn_destroy(h)


@pytest.mark.gpu
def test_snapshot_roundtrip_renders_identically(hn, root, tmp_path):
    path = str(tmp_path / "scene.ingp")
    blob, grid = make_snapshot(root, hn.synthetic, BaseInfo, path)
    a = hn.Testbed()
    a.reload_network_from_file(os.path.join(root, "configs", "nerf", "base.json"))
    a.set_params(blob)
    a.set_density_grid(grid, 1)
    a.snap_to_pixel_centers = True
    a.set_nerf_camera_matrix([[1, 0, 0, 0.5], [0, -1, 0, 0.5], [0, 0, -1, 2.25]])
    a.set_option("relative_focal_length", [1.2, 1.2])
    a.exposure = 0.5
    want = a.render(160, 120, 1, True)
    b = hn.Testbed()
    b.load_snapshot(path)                      # written by python msgpack + zlib (dynamic Huffman)
    b.snap_to_pixel_centers = True
    got = b.render(160, 120, 1, True)
    assert np.array_equal(got, want) and want[..., :3].max() > 0
    assert np.array_equal(b.get_occupancy_bitfield(), a.get_occupancy_bitfield())
    # our writer -> python's zlib/msgpack read it back -> our reader renders the same again
    out = str(tmp_path / "resaved.ingp")
    b.save_snapshot(out)
    assert open(out, "rb").read(2) == b"\x1f\x8b"   # the gzip container the reference writes
    back = msgpack.unpackb(gzip.decompress(open(out, "rb").read()), raw=False)
    assert back["snapshot"]["params_binary"] == blob.tobytes() and back["snapshot"]["version"] == 1
    assert back["encoding"]["n_levels"] == 8 and back["snapshot"]["camera"]["matrix"][2] == [0.0, 0.0, -1.0, 2.25]
    c = hn.Testbed()
    c.load_file(out)
    c.snap_to_pixel_centers = True
    assert np.array_equal(c.render(160, 120, 1, True), want)
    # hybrid scene json with a "Nerf" entry pointing at the snapshot (data/geometry/geometrypathsnerf.json format)
    import json
    obj = tmp_path / "tri.obj"
    obj.write_text("v 0.2 0.2 0.5\nv 0.8 0.2 0.5\nv 0.5 0.8 0.5\nf 1 2 3\n")
    scene = tmp_path / "geometrypathsnerf.json"
    scene.write_text(json.dumps({"geometry": [{"center": [0.5, 0.5, 0.5], "path": str(obj), "type": "Mesh"},
                                              {"center": [0.5, 0.5, 0.5], "path": out, "type": "Nerf"}]}))
    d = hn.Testbed()
    d.load_file(str(scene))
    img = d.render(96, 72, 1, True)
    assert d.frame_stats().n_samples > 0 and np.isfinite(img).all()
