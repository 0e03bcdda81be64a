"""Drop-in for the hot-path functions of `projects/nerf/utils/nerf_util.py` and the sampling loop of
`projects/neuralangelo/model.py:232-267`, running in libna_b200.so (csrc/render.cu)."""
import torch

from . import _lib
from ._lib import f32c, ptr

_L = _lib.lib
_unif_cache = {}


def _midpoint_unif(n_fine, device):
    """nerf_util.py:53-54: midpoints of linspace(0,1,Nf+1), cached per (Nf, device)."""
    key = (n_fine, str(device))
    if key not in _unif_cache:
        grid = torch.linspace(0, 1, n_fine + 1, device=device)
        _unif_cache[key] = (0.5 * (grid[:-1] + grid[1:])).contiguous()
    return _unif_cache[key]


def sample_dists(ray_size, dist_range, intvs, stratified, device="cuda"):
    """nerf_util.py:20-38 (pure index arithmetic + RNG: stays in torch).  Returns [B,R,N,1]."""
    batch_size, num_rays = ray_size
    dist_min, dist_max = dist_range
    if stratified:
        rands = torch.rand(batch_size, num_rays, intvs, 1, device=device)
    else:
        rands = torch.full((batch_size, num_rays, intvs, 1), 0.5, device=device)
    rands = rands + torch.arange(intvs, dtype=torch.float, device=device)[None, None, :, None]
    return rands / intvs * (dist_max - dist_min) + dist_min


@torch.no_grad()
def sample_dists_from_pdf(bin, weights, intvs_fine):
    """nerf_util.py:41-68.  bin [B,R,N,1] (the reference indexes bin[..., 0]), weights [B,R,N-1] -> [B,R,Nf,1]."""
    _lib.require_cuda(bin, weights)
    bins = f32c(bin[..., 0])
    w = f32c(weights)
    n = bins.shape[-1]
    rays = bins.numel() // n
    out = torch.empty(*bins.shape[:-1], intvs_fine, dtype=torch.float32, device=bins.device)
    _lib.check(_L().na_sample_pdf(ptr(bins), ptr(w), ptr(_midpoint_unif(intvs_fine, bins.device)), rays, n, intvs_fine,
                                  ptr(out), _lib.current_stream()), "na_sample_pdf")
    return out[..., None]


@torch.no_grad()
def sample_dists_hierarchical(dists, sdfs, inv_s, intvs_fine):
    """Model.sample_dists_hierarchical (model.py:250-267, robust=True).  dists, sdfs [B,R,N,1] -> [B,R,Nf,1]."""
    _lib.require_cuda(dists, sdfs)
    d, s = f32c(dists[..., 0]), f32c(sdfs[..., 0])
    n = d.shape[-1]
    rays = d.numel() // n
    out = torch.empty(*d.shape[:-1], intvs_fine, dtype=torch.float32, device=d.device)
    _lib.check(_L().na_neus_upsample(ptr(d), ptr(s), ptr(_midpoint_unif(intvs_fine, d.device)), float(inv_s), rays, n,
                                     intvs_fine, ptr(out), _lib.current_stream()), "na_neus_upsample")
    return out[..., None]


@torch.no_grad()
def merge_sorted(dists, sdfs, dists_fine, sdfs_fine=None):
    """cat + sort (+ gather of the SDF values) of model.py:241-247.  [B,R,N,1] ++ [B,R,Nf,1] -> [B,R,N+Nf,1]."""
    d, df = f32c(dists[..., 0]), f32c(dists_fine[..., 0])
    n, nf = d.shape[-1], df.shape[-1]
    rays = d.numel() // n
    carry = sdfs_fine is not None
    s = f32c(sdfs[..., 0]) if carry else None
    sf = f32c(sdfs_fine[..., 0]) if carry else None
    d_out = torch.empty(*d.shape[:-1], n + nf, dtype=torch.float32, device=d.device)
    s_out = torch.empty_like(d_out) if carry else None
    _lib.check(_L().na_merge_sorted(ptr(d), ptr(s), ptr(df), ptr(sf), rays, n, nf, ptr(d_out), ptr(s_out),
                                    _lib.current_stream()), "na_merge_sorted")
    return d_out[..., None], (s_out[..., None] if carry else None)


def intersect_with_sphere(center, ray_unit, radius=1.0):
    """nerf_util.py:199-205 (a dozen flops per ray: stays in torch).  NaN marks a miss (model.py:211)."""
    ctc = (center * center).sum(dim=-1, keepdim=True)
    ctv = (center * ray_unit).sum(dim=-1, keepdim=True)
    disc = (ctv ** 2 - (ctc - radius ** 2)).sqrt()
    return -ctv - disc, -ctv + disc
