"""Developer script: which ops launch the most kernels in one training step (eager, no graph)."""
import os, sys, runpy
os.environ["GRAPH"] = "0"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
# reuse the set-up of profile_train_step.py without its own profiling section
src = open(os.path.join(os.path.dirname(__file__), "profile_train_step.py")).read().split("from torch.profiler")[0]
ns = {"__file__": os.path.join(os.path.dirname(os.path.abspath(__file__)), "profile_train_step.py"), "__name__": "setup"}
exec(compile(src, "profile_train_step_setup", "exec"), ns)
ts, batch = ns["ts"], ns["batch"]
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    ts.step(batch(0))
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.device_type.name == "CUDA" or e.self_device_time_total > 0]
kern = [e for e in prof.key_averages() if e.self_device_time_total > 0]
kern.sort(key=lambda e: -e.count)
tot_n = sum(e.count for e in kern); tot_t = sum(e.self_device_time_total for e in kern)
print("kernel-launching ops per step:", tot_n, " total device us:", round(tot_t))
for e in kern[:45]:
    print(f"{e.count:5d} x {e.self_device_time_total / e.count:8.1f} us = {e.self_device_time_total:9.0f} us  {e.key[:90]}")
