"""Developer script: time the default-config training step and print a kernel-level breakdown."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuralangelo_b200 import build
from neuralangelo_b200.config import default_model_config, default_data_config
from neuralangelo_b200.model import Model
from neuralangelo_b200.trainer import TrainStep
build.build(verbose=False)
dev = torch.device("cuda:0")
torch.manual_seed(0)
cfg = default_model_config()
taps = int(os.environ.get("TAPS", "4"))
cfg.object.sdf.gradient.taps = taps
model = Model(cfg, default_data_config()).to(dev)
force16 = os.environ.get("FORCE16", "1") == "1"
ts = TrainStep(model, max_iter=500000, warm_up_end=5000, use_cuda_graph=os.environ.get("GRAPH", "0") == "1")
if force16:
    ts.current_iteration = 5000 + 16 * 5000   # all 16 levels active, eps = 1/2048

def look_at(eye):
    eye = torch.tensor(eye, dtype=torch.float32); fwd = -eye / eye.norm()
    right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd); right = right / right.norm()
    R = torch.stack([right, torch.linalg.cross(fwd, right), fwd]); return torch.cat([R, (-R @ eye)[:, None]], dim=1)
g = torch.Generator().manual_seed(0)
poses = []
for i in range(32):
    v = torch.randn(3, generator=g); v = v / v.norm() * 3.0; poses.append(look_at(v.tolist()))
intr = torch.tensor([[1111.0, 0, 400.0], [0, 1111.0, 400.0], [0, 0, 1.0]])
def batch(i):
    return dict(pose=poses[i % 32][None].to(dev), intr=intr[None].to(dev), idx=torch.tensor([i % 32], device=dev),
                ray_idx=torch.randint(0, 640000, (1, 512), generator=g).to(dev), image_sampled=torch.rand(1, 512, 3, generator=g).to(dev))
for i in range(5):
    ts.step(batch(i))
torch.cuda.synchronize()
t0 = time.perf_counter(); n = 20
for i in range(n):
    ts.step(batch(i))
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / n
print(f"taps={taps} force16={force16} active={model.neural_sdf.active_levels} step {dt*1e3:.2f} ms -> {512/dt:.0f} rays/s")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(3):
        ts.step(batch(i))
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))
