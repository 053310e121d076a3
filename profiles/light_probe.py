"""config 4 (3840x2160 hybrid frame, 1024x512 envmap, 16 light rays per mesh hit) on one GPU: frame time and the light pass's share;
under ncu (no warm-up frame: argv[1] = 0) the first launches are the light pass's own kernels:
  ncu --set full --clock-control none -k regex:"envlight|secondary_start|march_kernel|eval_tc|transmittance|compact" -c 10 -o gpurun_out/light python profiles/light_probe.py 0
options as name=value after the frame count."""
import os, sys, time, importlib
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 5
W, H = 3840, 2160
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
tb.set_envmap(hn.synthetic.envmap(1024, 512, seed=3))
tb.set_option("env_light_samples", 16.0)
tb.snap_to_pixel_centers = True
cam = hn.synthetic.default_camera(); tb.set_camera_cols(cam)
for kv in sys.argv[2:]:
    k, v = kv.split("="); tb.set_option(k, float(v))
frame = torch.zeros(H, W, 4, device="cuda"); depth = torch.zeros(H, W, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
if frames == 0:
    tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream); torch.cuda.synchronize(); sys.exit(0)
for _ in range(2): tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(frames): tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream)
torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / frames * 1e3
tb.set_option("profile", 1.0); tb.render_frame(cam, cam, W, H, frame.data_ptr(), depth.data_ptr(), 0, stream); st = tb.frame_stats()
print(f"config4 frame {ms:.2f} ms; profiled: envlight {st.ms_envlight:.2f} eval {st.ms_eval:.2f} march {st.ms_march:.2f} comp {st.ms_composite:.2f} mesh {st.ms_mesh:.2f} raygen {st.ms_raygen:.2f}; "
      f"hits {st.n_env_hits} light rays {st.n_light_rays} in volume {st.n_light_rays_volume} light samples {st.n_light_samples} ({st.n_light_samples / max(st.ms_envlight, 1e-9) / 1e6:.2f} Gsamples/s over the whole light pass)")
