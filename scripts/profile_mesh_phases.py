"""Developer script: where one block of the 2048^3 extraction spends its time (lattice build / encode / MLP / marching cubes)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuralangelo_b200 import build, mesh
from neuralangelo_b200.config import default_model_config
from neuralangelo_b200.modules import NeuralSDF
build.build(verbose=False)
dev = torch.device("cuda:0")
torch.manual_seed(0)
net = NeuralSDF(default_model_config().object.sdf).to(dev).eval()
net.warm_up_end = 5000
net.set_active_levels(500000)
bounds = np.array([[-1.0, 1.0]] * 3)
grid = mesh.LatticeGrid(bounds, intv=2.0 / 2048, block_res=128)
blocks = list(range(0, len(grid), len(grid) // 96))[:96]
ev = lambda: torch.cuda.Event(enable_timing=True)
tot = dict(lattice=0.0, encode=0.0, mlp=0.0, mc=0.0)
nv = 0
with torch.no_grad():
    for rep in range(2):
        for k in tot: tot[k] = 0.0
        t0 = time.perf_counter()
        for idx in blocks:
            e = [ev() for _ in range(5)]
            e[0].record()
            gx, gy, gz = (g.to(dev) for g in grid.block_axes(idx))
            xyz = torch.stack(torch.meshgrid(gx, gy, gz, indexing="ij"), dim=-1)
            e[1].record()
            planes = net.encode_planes(xyz.reshape(-1, 3))
            e[2].record()
            sdf = -net._packed_mlp().forward_planes(planes, planes.shape[1]).view(*xyz.shape[:-1])
            e[3].record()
            V, F, C = mesh.marching_cubes(sdf, xyz, 2.0 / 2048, None, False)
            e[4].record()
            torch.cuda.synchronize()
            nv += V.shape[0]
            for k, i in zip(tot, range(4)):
                tot[k] += e[i].elapsed_time(e[i + 1])
        wall = (time.perf_counter() - t0) * 1e3
print({k: round(v / len(blocks), 3) for k, v in tot.items()}, "ms per block; wall per block", round(wall / len(blocks), 3), "ms; vertices", nv)
