"""CPU restatement of the reference's NeuS rendering / resampling math -- TEST INFRASTRUCTURE ONLY.

Pinned: every function here is checked against outputs of the reference's own code (imported from
/root/reference in the authoring container by oracle/make_golden.py; fixtures in tests/golden/
render_*.npz) and against the known-answer table of SURVEY.md section 4, in tests/test_oracle_render.py.

Each function cites the reference lines it restates.  Inputs/outputs are torch tensors (any float
dtype; tests use float32 for parity and float64 for gradient checks) laid out [rays, samples(, k)].
Gradients come from torch autograd over these formulas.
"""
import torch


def _append_far(dists, dist_far):
    """dists [R,N], dist_far [R] or None -> interval lengths [R,N] (last one closed by dist_far / 1e10).
    render.py:79-82, model.py:283-286."""
    if dist_far is None:
        far = torch.full_like(dists[:, :1], 1e10)
    else:
        far = dist_far.reshape(-1, 1).to(dists.dtype)
    ext = torch.cat([dists, far], dim=1)
    return ext[:, 1:] - ext[:, :-1]


def density_alphas(densities, dists, dist_far=None):
    """render.py:68-84 volume_rendering_alphas_dist: alpha = 1 - exp(-sigma * delta)."""
    return 1.0 - torch.exp(-densities * _append_far(dists, dist_far))


def alpha_weights(alphas):
    """render.py:87-99 alpha_compositing_weights: w_i = alpha_i * prod_{j<i} (1 - alpha_j)."""
    one_minus = 1.0 - alphas
    shifted = torch.cat([torch.ones_like(one_minus[:, :1]), one_minus[:, :-1]], dim=1)
    return alphas * torch.cumprod(shifted, dim=1)


def composite(quantities, weights):
    """render.py:102-112: sum_i q_i * w_i.  quantities [R,N,K], weights [R,N] -> [R,K]."""
    return (quantities * weights[..., None]).sum(dim=1)


def iter_cos(true_cos, anneal_ratio):
    """model.py:294-298 _get_iter_cos with anneal_ratio = min(progress / anneal_end, 1)."""
    early = torch.relu(0.5 - 0.5 * true_cos)
    late = torch.relu(-true_cos)
    return -(early * (1.0 - anneal_ratio) + late * anneal_ratio)


def _section_alpha(center_sdf, slope, delta, inv_s, eps=1e-5):
    """Shared by model.py:287-290 and model.py:261-265: opacity of one section from the SDF
    estimated at its two ends through a logistic CDF of sharpness inv_s."""
    front = torch.sigmoid((center_sdf - 0.5 * slope * delta) * inv_s)
    back = torch.sigmoid((center_sdf + 0.5 * slope * delta) * inv_s)
    return ((front - back) / (front + eps)).clamp(0.0, 1.0)


def neus_alphas(sdf, grad, dists, ray_unit, dist_far, s_var, anneal_ratio):
    """model.py:275-292 compute_neus_alphas.  sdf [R,N], grad [R,N,3], dists [R,N], ray_unit [R,3],
    dist_far [R] or None, s_var scalar tensor."""
    cos = (ray_unit[:, None, :] * grad).sum(-1)
    return _section_alpha(sdf, iter_cos(cos, anneal_ratio), _append_far(dists, dist_far), torch.exp(s_var))


def sample_pdf(bins, weights, unif):
    """nerf_util.py:41-68 sample_dists_from_pdf.  bins [R,N], weights [R,N-1], unif [Nf] (the midpoints
    of linspace(0,1,Nf+1), passed in so that oracle and kernel see identical u values)."""
    pdf = weights / weights.abs().sum(-1, keepdim=True).clamp_min(1e-12)
    cdf = torch.cat([torch.zeros_like(pdf[:, :1]), torch.cumsum(pdf, dim=-1)], dim=-1)  # [R,N]
    u = unif[None, :].expand(cdf.shape[0], -1).contiguous()
    first_above = torch.searchsorted(cdf, u, right=True)
    lo = (first_above - 1).clamp(min=0)
    hi = first_above.clamp(max=cdf.shape[-1] - 1)
    c_lo, c_hi = cdf.gather(1, lo), cdf.gather(1, hi)
    b_lo, b_hi = bins.gather(1, lo), bins.gather(1, hi)
    frac = (u - c_lo) / (c_hi - c_lo + 1e-8)
    return b_lo + frac * (b_hi - b_lo)


def midpoint_unif(n_fine, dtype=torch.float32):
    """nerf_util.py:53-54: 0.5 * (grid[:-1] + grid[1:]) with grid = linspace(0, 1, Nf+1)."""
    grid = torch.linspace(0, 1, n_fine + 1, dtype=dtype)
    return 0.5 * (grid[:-1] + grid[1:])


def neus_upsample(dists, sdfs, unif, inv_s):
    """model.py:250-267 sample_dists_hierarchical (robust=True).  dists, sdfs [R,N] -> [R,Nf]."""
    delta = dists[:, 1:] - dists[:, :-1]
    mid = 0.5 * (sdfs[:, 1:] + sdfs[:, :-1])
    slope = (sdfs[:, 1:] - sdfs[:, :-1]) / (delta + 1e-5)
    before = torch.cat([torch.zeros_like(slope[:, :1]), slope[:, :-1]], dim=1)
    slope = torch.minimum(before, slope)
    weights = alpha_weights(_section_alpha(mid, slope, delta, inv_s))
    return sample_pdf(dists, weights, unif)


def merge_sorted(dists, sdfs, dists_fine, sdfs_fine):
    """model.py:241-247: cat, sort along samples, and carry the SDF values along."""
    both = torch.cat([dists, dists_fine], dim=1)
    merged, order = both.sort(dim=1, stable=True)
    if sdfs is None:
        return merged, None
    return merged, torch.cat([sdfs, sdfs_fine], dim=1).gather(1, order)


def neus_render(sdf, grad, dists_obj, ray_unit, dist_far, s_var, anneal_ratio, rgb_obj, dens_bg=None,
                dists_bg=None, rgb_bg=None):
    """model.py:124-141 (+ :168-169, :198): object NeuS alphas ++ background density alphas ->
    weights -> composited RGB.  Returns (rgb [R,3], weights [R,No+Nb], opacity [R])."""
    alphas = neus_alphas(sdf, grad, dists_obj, ray_unit, dist_far, s_var, anneal_ratio)
    colors = rgb_obj
    if dens_bg is not None:
        alphas = torch.cat([alphas, density_alphas(dens_bg, dists_bg, None)], dim=1)
        colors = torch.cat([rgb_obj, rgb_bg], dim=1)
    weights = alpha_weights(alphas)
    return composite(colors, weights), weights, weights.sum(-1)
