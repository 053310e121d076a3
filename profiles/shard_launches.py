"""rank 0's share of the benched 1080p hybrid frame at world = int(argv[1]) on ONE GPU, for an ncu launch list
(ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv python profiles/shard_launches.py 8): true per-kernel
durations of a shard, without the event/sync brackets of the library's profile mode. Prints the un-profiled frame time too."""
import os, sys, time, importlib
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 20
W, H = 1920, 1080
tb = hn.Testbed(device=0)
for kv in sys.argv[(3 if "shard" in __file__ else 2):]:   # options that act at scene-build time (bvh_sah) must be set before build_scene
    k, v = kv.split("="); tb.set_option(k, float(v))
tb.reload_network_from_file(os.path.join(ROOT, "configs", "nerf", "base.json"))
info = tb.network_info()
tb.set_params(hn.synthetic.random_params(info, 1337, 1e-4))
tb.set_density_grid(hn.synthetic.sphere_density_grid(0.35), 1)
tb.add_mesh(hn.synthetic.torus_mesh(1024, 512), normalize=True); tb.build_scene()
tb.set_option("render_aabb", tb.get_option("mesh_aabb", 6))
tb.set_option("compat_nearest_aabb_only", 1.0)
tb.snap_to_pixel_centers = True
cam = hn.synthetic.default_camera(); tb.set_camera_cols(cam)
for kv in sys.argv[3:]:
    k, v = kv.split("="); tb.set_option(k, float(v))
frame = torch.zeros(H, W, 4, device="cuda"); depth = torch.zeros(H, W, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
tb.set_tile_shard(0, world, 64, 32)
for _ in range(3): tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(frames): tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / frames * 1e3
print(f"world {world}: rank-0 shard frame {ms:.3f} ms (wall clock, {frames} frames, no gather)")
