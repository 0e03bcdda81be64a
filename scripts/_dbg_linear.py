import os, sys
os.environ["GRAPH"] = "0"
sys.path.insert(0, "/root/repo"); sys.path.insert(0, ".")
import torch
src = open("scripts/profile_train_step.py").read().split("from torch.profiler")[0].split("for i in range(5):")[0]
ns = {"__file__": os.path.abspath("scripts/profile_train_step.py"), "__name__": "setup"}
exec(compile(src, "setup", "exec"), ns)
ts, batch = ns["ts"], ns["batch"]
from neuralangelo_b200 import mlp
orig = mlp._LinearTallRows.forward
def fwd(ctx, x, w, b):
    print("linear", tuple(x.shape), x.stride(), x.data_ptr() % 16, "| w", tuple(w.shape), w.stride(), w.data_ptr() % 16, w.is_contiguous(), "| b", b.data_ptr() % 16, b.stride())
    return orig(ctx, x, w, b)
mlp._LinearTallRows.forward = staticmethod(fwd)
ts.step(batch(0))
torch.cuda.synchronize()
