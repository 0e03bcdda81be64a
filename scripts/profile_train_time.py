"""Developer script: device time of one training step (eager, no graph) grouped by kernel, sorted by total time."""
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
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    ts.step(batch(5))
    torch.cuda.synchronize()
kern = [e for e in prof.key_averages() if e.self_device_time_total > 0]
kern.sort(key=lambda e: -e.self_device_time_total)
tot_n = sum(e.count for e in kern); tot_t = sum(e.self_device_time_total for e in kern)
print("kernel launches per step:", tot_n, " total device us:", round(tot_t))
small = [e for e in kern if e.self_device_time_total / e.count < 6.0]
print("launches under 6 us:", sum(e.count for e in small), " their total us:", round(sum(e.self_device_time_total for e in small)))
for e in kern[:40]:
    print(f"{e.count:5d} x {e.self_device_time_total / e.count:8.1f} us = {e.self_device_time_total:9.0f} us  {e.key[:100]}")
