"""One weight-gradient launch of the largest shape of a training step (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuralangelo_b200 import build
from neuralangelo_b200.mlp import linear_wgrad
build.build(verbose=False)
dev = torch.device("cuda:0")
dy = torch.randn(262144, 256, device=dev) * 1e-6
x = torch.relu(torch.randn(262144, 131, device=dev))
for _ in range(3):
    linear_wgrad(dy, x)
torch.cuda.synchronize()
print("ok")
