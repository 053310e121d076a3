"""config3 1080p frame on one GPU: separate compositing kernel vs compositing fused into the tcgen05 network kernel's epilogue."""
import os, sys, importlib
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
base=None
for fuse in (0, 1, 0, 1):
    tb.set_option("fuse_composite", float(fuse))
    for _ in range(3): tb.render_frame(cam,cam,W,H,frame.data_ptr(),depth.data_ptr(),0,stream)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): tb.render_frame(cam,cam,W,H,frame.data_ptr(),depth.data_ptr(),0,stream)
    e1.record(); torch.cuda.synchronize(); ms=e0.elapsed_time(e1)/20
    img=frame.cpu().numpy().copy()
    if base is None: base=img
    tb.set_option("profile",1.0)
    tb.render_frame(cam,cam,W,H,frame.data_ptr(),depth.data_ptr(),0,stream); st=tb.frame_stats()
    tb.set_option("profile",0.0)
    print(f"fuse_composite {fuse}: frame {ms:.3f} ms  eval {st.ms_eval:.3f} march {st.ms_march:.3f} comp {st.ms_composite:.3f} mesh {st.ms_mesh:.3f} raygen {st.ms_raygen:.3f}  max|diff| {np.abs(img-base).max():.2e}")
