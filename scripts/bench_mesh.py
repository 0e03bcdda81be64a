"""Developer script: extract_mesh timing with the default hash-grid model (random init) and an analytic sphere."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuralangelo_b200 import build, mesh
from neuralangelo_b200.config import default_model_config
from neuralangelo_b200.modules import NeuralSDF
build.build(verbose=False)
dev = torch.device("cuda:0")
torch.manual_seed(0)
res = int(os.environ.get("RES", "512"))
cfg = default_model_config()
sdf_net = NeuralSDF(cfg.object.sdf).to(dev).eval()
sdf_net.warm_up_end = 5000
sdf_net.set_active_levels(500000)
bounds = np.array([[-1.0, 1.0]] * 3)
sdf_func = lambda x: -sdf_net.sdf(x)
for name, fn in (("hashgrid_model", sdf_func), ("analytic_sphere", lambda x: -(x.norm(dim=-1, keepdim=True) - 0.5131))):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    m = mesh.extract_mesh(fn, bounds, intv=2.0 / res, block_res=128)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    g = mesh.LatticeGrid(bounds, 2.0 / res, 128)
    npts = sum(int(np.prod([len(a) for a in g.block_axes(i)])) for i in range(len(g)))
    print(f"{name}: res {res}: {len(g)} blocks, {npts/1e6:.1f} M lattice points, {dt:.3f} s -> {npts/dt/1e6:.1f} M pts/s; V={len(m.vertices)} F={len(m.faces)}")
