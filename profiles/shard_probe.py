"""Per-rank frame time of the tile-sharded render path measured on ONE GPU (rank 0 of world = 1, 2, 4, 8): separates the
scaling of the kernels themselves from the NCCL gather. Output of the last run: profiles/r01_shard_probe.txt"""
import os, sys, time, importlib
import numpy as np, torch
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
W,H=1920,1080
tb = hn.Testbed(device=0)
tb.reload_network_from_file(os.path.join(ROOT,"configs","nerf","base.json"))
info=tb.network_info()
tb.set_params(hn.synthetic.random_params(info,1337,1e-4))
tb.set_density_grid(hn.synthetic.sphere_density_grid(0.35),1)
tb.add_mesh(hn.synthetic.torus_mesh(1024,512)); tb.build_scene()
tb.snap_to_pixel_centers=True
cam=hn.synthetic.default_camera(); tb.set_camera_cols(cam)
frame=torch.zeros(H,W,4,device="cuda"); depth=torch.zeros(H,W,device="cuda")
stream=torch.cuda.current_stream().cuda_stream
for world in (1,2,4,8):
    tb.set_tile_shard(0, world, 64, 32)
    for _ in range(3): tb.render_frame(cam,cam,W,H,frame.data_ptr(),depth.data_ptr(),0,stream)
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(20): tb.render_frame(cam,cam,W,H,frame.data_ptr(),depth.data_ptr(),0,stream)
    torch.cuda.synchronize(); ms=(time.perf_counter()-t0)/20*1e3
    tb.set_option("profile",1.0); tb.render_frame(cam,cam,W,H,frame.data_ptr(),depth.data_ptr(),0,stream); st=tb.frame_stats(); tb.set_option("profile",0.0)
    print(f"world {world}: rank0 frame {ms:.3f} ms; kernels eval {st.ms_eval:.3f} march {st.ms_march:.3f} comp {st.ms_composite:.3f} mesh {st.ms_mesh:.3f} raygen {st.ms_raygen:.3f} sum {st.ms_eval+st.ms_march+st.ms_composite+st.ms_mesh+st.ms_raygen:.3f} passes {st.n_passes}")
