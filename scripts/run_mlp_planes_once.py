"""Developer script for profiler captures: a few launches of the pipelined SDF-only MLP (chunk planes, 2^21 rows) and of
the two-layer forward."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuralangelo_b200 import build
build.build(verbose=False)
from neuralangelo_b200.mlp import MLPforNeuralSDF, get_activation
from neuralangelo_b200.fused_mlp import PackedSdfMlp
dev = torch.device("cuda:0")
torch.manual_seed(0)
mlp = MLPforNeuralSDF([131, 256, 256], activ=get_activation("softplus", beta=100.0), use_weightnorm=True, geometric_init=True, out_bias=0.5).to(dev)
packed = PackedSdfMlp(mlp, 100.0)
m = 1 << 21
planes = torch.randn(17, m, 8, device=dev)
rows = torch.randn(m, 131, device=dev)
for _ in range(3):
    packed.forward_planes(planes, m)
    packed.forward(rows, with_sdf=True, with_feat=True)
torch.cuda.synchronize()
print("done")
