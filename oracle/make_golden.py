"""Generates tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference
through oracle/reference_shim.py) on seeded inputs.  Run in the authoring container only:

    python -m oracle.make_golden

The fixtures are small (a few hundred KB) and committed; tests/test_oracle_render.py checks
oracle/render_ref.py against them, so the oracle stays pinned on machines without the reference.
"""
import os
import types

import numpy as np
import torch

from . import reference_shim

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy()


def render_fixtures():
    reference_shim.install()
    from projects.nerf.utils import render, nerf_util
    from projects.neuralangelo.model import Model

    g = torch.Generator().manual_seed(20231106)
    B, R, N, NB, NF = 2, 24, 37, 9, 16
    out = {}
    # ---- render.py functions ------------------------------------------------------------
    alphas = torch.rand(B, R, N, generator=g) ** 3
    alphas[0, 0, 5] = 1.0   # exact-one alpha: transmittance hits 0 (clip at model.py:290 makes this reachable)
    alphas[0, 1, :] = 0.0
    out["aw_alphas"] = _np(alphas)
    out["aw_weights"] = _np(render.alpha_compositing_weights(alphas))
    q = torch.randn(B, R, N, 3, generator=g)
    out["comp_q"] = _np(q)
    out["comp_out"] = _np(render.composite(q, render.alpha_compositing_weights(alphas)))
    dens = torch.rand(B, R, NB, generator=g) * 5
    dists = torch.sort(torch.rand(B, R, NB, 1, generator=g) * 4 + 0.5, dim=2).values
    far = dists[..., -1:, :] + torch.rand(B, R, 1, 1, generator=g)
    out["da_dens"], out["da_dists"], out["da_far"] = _np(dens), _np(dists), _np(far)
    out["da_alphas_far"] = _np(render.volume_rendering_alphas_dist(dens, dists, far))
    out["da_alphas_nofar"] = _np(render.volume_rendering_alphas_dist(dens, dists, None))
    # ---- nerf_util.sample_dists_from_pdf -------------------------------------------------
    bins = torch.sort(torch.rand(B, R, N, 1, generator=g) * 3 + 1, dim=2).values
    w = torch.rand(B, R, N - 1, generator=g) ** 4
    w[0, 2, 10:20] = 0.0    # flat CDF stretch
    w[1, 3, :] = 0.0        # all-zero weights
    out["pdf_bins"], out["pdf_w"] = _np(bins), _np(w)
    out["pdf_out"] = _np(nerf_util.sample_dists_from_pdf(bins, w, NF))
    # ---- Model methods (unbound: they only touch self.s_var / self.anneal_end / self.sample_dists_from_pdf)
    stub = types.SimpleNamespace()
    stub.s_var = torch.tensor(3.0)
    stub.anneal_end = 0.1
    stub._get_iter_cos = lambda c, progress=1.: Model._get_iter_cos(stub, c, progress=progress)
    import functools
    stub.sample_dists_from_pdf = functools.partial(nerf_util.sample_dists_from_pdf, intvs_fine=NF)
    ray = torch.nn.functional.normalize(torch.randn(B, R, 3, generator=g), dim=-1)
    sd = torch.sort(torch.rand(B, R, N, 1, generator=g) * 2 + 1, dim=2).values
    # SDF crossing zero along the ray plus noise; gradients roughly anti-parallel to the ray
    sdfs = (2.0 - sd) * 0.4 + 0.02 * torch.randn(B, R, N, 1, generator=g)
    grads = -ray[:, :, None, :] * (0.8 + 0.4 * torch.rand(B, R, N, 1, generator=g)) + \
        0.3 * torch.randn(B, R, N, 3, generator=g)
    dfar = sd[..., -1:, :] + 0.05 + torch.rand(B, R, 1, 1, generator=g) * 0.1
    out["na_ray"], out["na_sdf"], out["na_grad"], out["na_dists"], out["na_far"] = map(_np, (ray, sdfs, grads, sd, dfar))
    for tag, prog in (("p0", 0.0), ("p05", 0.05), ("p1", 1.0)):
        out[f"na_alphas_{tag}"] = _np(Model.compute_neus_alphas(stub, ray, sdfs, grads, sd, dist_far=dfar, progress=prog))
    out["na_alphas_nofar"] = _np(Model.compute_neus_alphas(stub, ray, sdfs, grads, sd, dist_far=None, progress=1.0))
    # gradients of a scalar loss through compute_neus_alphas -> weights -> composite
    sdfs_g, grads_g = sdfs.clone().requires_grad_(True), grads.clone().requires_grad_(True)
    stub.s_var = torch.tensor(3.0, requires_grad=True)
    rgbs = torch.rand(B, R, N, 3, generator=g)
    a = Model.compute_neus_alphas(stub, ray, sdfs_g, grads_g, sd, dist_far=dfar, progress=0.05)
    wts = render.alpha_compositing_weights(a)
    rgb = render.composite(rgbs, wts)
    tgt = torch.rand(B, R, 3, generator=g)
    loss = ((rgb - tgt) ** 2).sum()
    loss.backward()
    out["bw_rgbs"], out["bw_tgt"], out["bw_rgb"] = _np(rgbs), _np(tgt), _np(rgb)
    out["bw_d_sdf"], out["bw_d_grad"], out["bw_d_svar"] = _np(sdfs_g.grad), _np(grads_g.grad), _np(stub.s_var.grad)
    stub.s_var = torch.tensor(3.0)
    # hierarchical upsampling
    for h in (0, 2):
        out[f"up_fine_h{h}"] = _np(Model.sample_dists_hierarchical(stub, sd, sdfs, inv_s=64 * 2 ** h))
    # known answers of SURVEY.md section 4
    kat_s = torch.tensor([.3, .1, -.1, -.3]).view(1, 1, 4, 1)
    kat_d = torch.tensor([1, 1.2, 1.4, 1.6]).view(1, 1, 4, 1)
    NF4 = functools.partial(nerf_util.sample_dists_from_pdf, intvs_fine=4)
    stub.sample_dists_from_pdf = NF4
    out["kat_up"] = _np(Model.sample_dists_hierarchical(stub, kat_d, kat_s, inv_s=64))
    np.savez_compressed(os.path.join(OUT, "render_golden.npz"), **out)
    print("wrote render_golden.npz with", len(out), "arrays")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    render_fixtures()
