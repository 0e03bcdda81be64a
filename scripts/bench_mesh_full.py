"""Full-size isosurface extraction (BASELINE.json configs[4]): default hash-grid model, resolution 2048, block 128,
blocks sharded across the ranks of the launch (1 process = whole job).  Prints seconds end to end on rank 0."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from neuralangelo_b200 import build, mesh, parallel
from neuralangelo_b200.config import default_model_config
from neuralangelo_b200.modules import NeuralSDF
build.build(verbose=False)
rank, local, world = parallel.init_dist()
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
torch.manual_seed(0)
res = int(os.environ.get("RES", "2048"))
net = NeuralSDF(default_model_config().object.sdf).to(dev).eval()
net.warm_up_end = 5000
net.set_active_levels(500000)
bounds = np.array([[-1.0, 1.0]] * 3)
fn = mesh.LatticeSDF(net)   # block queries straight from the lattice axes (no xyz tensor)
mesh.extract_mesh(fn, bounds, intv=2.0 / 128, block_res=128)
if world > 1: dist.barrier()
torch.cuda.synchronize(); t0 = time.perf_counter()
stats = {}
m = mesh.extract_mesh(fn, bounds, intv=2.0 / res, block_res=128, stats=stats)
print('rank', rank, {k: round(v, 3) for k, v in stats.items()}, flush=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
if world > 1:
    t = torch.tensor([dt], device=dev, dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t)
if rank == 0:
    g = mesh.LatticeGrid(bounds, 2.0 / res, 128)
    print({"resolution": res, "n_gpus": world, "blocks": len(g), "seconds": round(dt, 3), "vertices": len(m.vertices), "faces": len(m.faces),
           "peak_mem_GB": round(torch.cuda.max_memory_allocated() / 1e9, 2)})
if world > 1: dist.destroy_process_group()
