"""Training step (CUDA graph, taps 4, 16 levels) with the colour / background MLPs on library GEMMs vs on the chain kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import torch
import bench
from neuralangelo_b200 import modules

out = {}
for flag in (False, True):
    modules.FUSED_CHAIN_IN_TRAINING = flag
    r = bench.extra_train_step(0, 1, torch.device("cuda:0"), steps=100, warmup=30, taps=4)
    out["chain" if flag else "library"] = {k: r[k] for k in ("ms_per_step", "phases_ms", "last_loss")}
print(json.dumps(out))
