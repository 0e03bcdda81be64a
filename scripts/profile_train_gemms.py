"""Developer script: the library GEMMs of one training step with their shapes and kernels."""
import os, sys
os.environ["GRAPH"] = "0"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
src = open(os.path.join(os.path.dirname(__file__), "profile_train_step.py")).read().split("from torch.profiler")[0]
ns = {"__file__": os.path.join(os.path.dirname(os.path.abspath(__file__)), "profile_train_step.py"), "__name__": "setup"}
exec(compile(src, "profile_train_step_setup", "exec"), ns)
ts, batch = ns["ts"], ns["batch"]
for i in range(3):
    ts.step(batch(i))
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
    ts.step(batch(5))
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.name in ("aten::addmm", "aten::mm", "aten::bmm", "aten::matmul", "aten::linear")]
for e in evs:
    if e.name in ("aten::addmm", "aten::mm", "aten::bmm"):
        print(e.name, e.input_shapes, "device us", round(e.device_time_total, 1))
kern = {}
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA and ("gemm" in e.name or "cutlass" in e.name or "sgemm" in e.name):
        kern.setdefault(e.name[:110], []).append(e.device_time_total)
for k, v in sorted(kern.items(), key=lambda kv: -sum(kv[1])):
    print(len(v), "x", round(sum(v) / len(v), 1), "us", k)
