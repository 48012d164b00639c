"""Where does a mapping-window step spend its time?  torch.profiler over a few steps (dev tool)."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from monogs_dev_b200 import synthetic as syn, mapping
from torch.profiler import profile, ProfilerActivity

dev = torch.device("cuda", 0)
sc = syn.make_scene(bench.WORKLOAD, device=dev, n_views=bench.N_VIEWS)
kfs, host = bench.build_window(dev, mapping.render, mapping.GaussianModel, None, sc)
gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation)
win = mapping.MappingWindow(gm, kfs, sc.bg)
for _ in range(3): win.step()
torch.cuda.synchronize()
t = time.time()
for _ in range(5): win.step()
torch.cuda.synchronize()
print("wall ms/step", (time.time() - t) / 5 * 1e3)
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3): win.step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=25, max_name_column_width=60))
