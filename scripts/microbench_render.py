"""Compositing microbenchmark: fused NeuS renderer forward + backward on 2^20 rays x
(128 object + 32 background) samples; reports achieved GB/s of algorithmic bytes against the HBM peak."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuralangelo_b200 import build, render
build.build(verbose=False)
dev = torch.device("cuda:0")
R, N, NB = 1 << int(os.environ.get("LOG2_RAYS", "20")), 128, 32
g = torch.Generator(device=dev).manual_seed(0)
ray = torch.nn.functional.normalize(torch.randn(1, R, 3, device=dev, generator=g), dim=-1)
dists = torch.sort(torch.rand(1, R, N, 1, device=dev, generator=g) * 2 + 1, dim=2).values
sdf = ((2.0 - dists) * 0.4 + 0.02 * torch.randn(1, R, N, 1, device=dev, generator=g)).requires_grad_(True)
grad = (-ray[:, :, None, :] + 0.3 * torch.randn(1, R, N, 3, device=dev, generator=g)).requires_grad_(True)
rgb = torch.rand(1, R, N, 3, device=dev, generator=g).requires_grad_(True)
dens = torch.rand(1, R, NB, device=dev, generator=g).requires_grad_(True)
dbg = torch.sort(torch.rand(1, R, NB, 1, device=dev, generator=g) * 10 + 3.2, dim=2).values
rgbb = torch.rand(1, R, NB, 3, device=dev, generator=g).requires_grad_(True)
far = dists[:, :, -1:, :] + 0.1
s_var = torch.tensor(3.0, device=dev, requires_grad=True)
ITERS, WARM = int(os.environ.get("NA_MB_ITERS", "5")), int(os.environ.get("NA_MB_WARM", "2"))
def timeit(fn):
    for _ in range(WARM): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(ITERS):
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(t))
    return float(np.median(ts))
def fwd(want_weights):
    with torch.no_grad():
        return render.neus_render(ray, sdf, grad, dists, s_var, rgb, dist_far=far, densities_bg=dens, dists_bg=dbg, rgbs_bg=rgbb, want_weights=want_weights)
# algorithmic bytes per ray (forward): object sample sdf 4 + grad 12 + dist 4 + rgb 12; bg sample dens 4 + dist 4 + rgb 12; + weights 4/sample (optional) + rgb/opacity out
b_in = N * 32 + NB * 20 + 12 + 4
res = {}
for ww in (False, True):
    ms = timeit(lambda: fwd(ww))
    by = R * (b_in + 16 + (160 * 4 if ww else 0))
    res["fwd_weights" if ww else "fwd"] = dict(ms=ms, GB=by / 1e9, GBps=by / ms / 1e6)
out, _, _ = render.neus_render(ray, sdf, grad, dists, s_var, rgb, dist_far=far, densities_bg=dens, dists_bg=dbg, rgbs_bg=rgbb, want_weights=False)
d_rgb = torch.randn_like(out)
def bwd():
    torch.autograd.grad(out, [sdf, grad, rgb, dens, rgbb, s_var], d_rgb, retain_graph=True)
ms = timeit(bwd)
by = R * (b_in + 12 + N * (4 + 12 + 12) + NB * (4 + 12))   # re-read inputs + d_rgb, write d_sdf, d_grad, d_rgb_obj, d_dens, d_rgb_bg
res["bwd"] = dict(ms=ms, GB=by / 1e9, GBps=by / ms / 1e6)
print(json.dumps(res, indent=1))
