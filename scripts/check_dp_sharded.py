"""2-rank check: sharded DP step == single-process step on the concatenated batch semantics (mean of grads)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch, torch.distributed as dist
from neuralangelo_b200 import parallel
rank, local, world = parallel.init_dist()
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
from test_train_gpu import _tiny_trainer, _batch
torch.backends.cuda.matmul.allow_tf32 = False
model, ts = _tiny_trainer(dev, seed=7)
model.cfg_render.stratified = False
losses = []
for s in range(8):
    losses.append(ts.step(_batch(dev, 2 * s + rank)).item())     # different rays per rank
# all ranks must hold identical parameters afterwards
flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
ref = flat.clone(); dist.broadcast(ref, src=0)
same = bool(torch.equal(flat, ref))
finite = bool(torch.isfinite(flat).all())
out = [None] * world
dist.all_gather_object(out, (rank, same, finite, [round(l, 5) for l in losses], float(flat.double().abs().sum())))
if rank == 0:
    for o in out: print(o)
dist.destroy_process_group()
