"""Cost of one occupancy-grid update (update_density_grid_nerf) and of training steps with / without the automatic update."""
import os, sys, time, importlib
import numpy as np, torch
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
hn = importlib.import_module("surface-irradiance-estimation-from-neural-radiance-fields_b200")
tb = hn.Testbed(device=0)
tb.reload_network_from_file(os.path.join(ROOT,"configs","nerf","base.json"))
info=tb.network_info()
tb.set_params(hn.synthetic.random_params(info,1337,1e-4))
tb.set_density_grid(hn.synthetic.sphere_density_grid(0.35),1)
imgs,cams,focal=hn.synthetic.sphere_dataset(64,256)
tb.set_training_data(imgs,cams,focal)
tb.reset_trainer(1337)
for _ in range(40): tb.train(1)
torch.cuda.synchronize()
def timeit(f,n):
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/n*1e3
print("train step, auto grid on : %.3f ms" % timeit(lambda: tb.train(1), 64))
tb.set_option("train_auto_grid", 0.0)
print("train step, auto grid off: %.3f ms" % timeit(lambda: tb.train(1), 64))
M=128**3
print("update_density_grid(M/4 uniform + M/4 nonuniform): %.3f ms" % timeit(lambda: tb.update_density_grid(0.95, M//4, M//4), 8))
print("update_density_grid(M uniform, warm-up form)      : %.3f ms" % timeit(lambda: tb.update_density_grid(0.95, M, 0), 8))
