"""GPU parity of the element-wise glue kernels (csrc/glue.cu) against the reference's own tensor expressions evaluated
with torch in float64 on the CPU: values and gradients.  Tolerance: fp32 arithmetic of the same formulas -> 2e-6
relative to the tensor scale forward, 2e-5 on gradients."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _close(got, ref, tol, name=""):
    got, ref = got.detach().cpu().double(), ref.detach().double()
    assert got.shape == ref.shape, (name, got.shape, ref.shape)
    scale = max(ref.abs().max().item(), 1e-30)
    assert (got - ref).abs().max().item() <= tol * scale, (name, (got - ref).abs().max().item(), scale)


@pytest.mark.parametrize("levels,n_app", [(3, 0), (4, 8), (1, 0)])
def test_rgb_input_matches_reference_expression(na_lib, cuda_device, levels, n_app):
    """cat([points, SH(rays), normalize(gradients), feats, app]) of modules.py:209-229 + model.py:181."""
    from neuralangelo_b200 import glue
    from neuralangelo_b200.spherical_harmonics import get_spherical_harmonics
    g = torch.Generator().manual_seed(levels)
    B, R, N, F = 2, 37, 5, 64
    pts = torch.randn(B, R, N, 3, generator=g)
    ray = torch.nn.functional.normalize(torch.randn(B, R, 3, generator=g), dim=-1)
    grad = torch.randn(B, R, N, 3, generator=g) * 3
    grad[0, 0, 0] = 0.0                      # a zero gradient: normalize's clamp branch
    feat = torch.randn(B, R, N, F, generator=g)
    app = torch.randn(B, R, N, n_app, generator=g) if n_app else None
    w = torch.randn(B, R, N, 6 + (levels + 1) ** 2 + F + n_app, generator=g)
    # reference (float64)
    gd, fd = grad.double().requires_grad_(True), feat.double().requires_grad_(True)
    rays = ray.double()[..., None, :].expand(B, R, N, 3)
    parts = [pts.double(), get_spherical_harmonics(rays, levels), torch.nn.functional.normalize(gd, dim=-1), fd]
    if app is not None:
        parts.append(app.double())
    ref = torch.cat(parts, dim=-1)
    (ref * w.double()).sum().backward()
    d = cuda_device
    gk, fk = grad.to(d).requires_grad_(True), feat.to(d).requires_grad_(True)
    out = glue.rgb_input(pts.to(d), ray.to(d), gk, fk, None if app is None else app.to(d), levels)
    _close(out, ref, 2e-6, "rgb_input")
    (out * w.to(d)).sum().backward()
    _close(fk.grad, fd.grad, 1e-6, "d_feat")
    mask = torch.ones(B, R, N, dtype=torch.bool)
    mask[0, 0, 0] = False                    # at g = 0 torch's normalize backward and the clamp branch differ by design
    _close(gk.grad.cpu()[mask], gd.grad[mask], 2e-5, "d_grad")


@pytest.mark.parametrize("taps,training", [(6, True), (4, True), (6, False), (4, False)])
def test_tap_combine_matches_reference_expression(na_lib, cuda_device, taps, training):
    """modules.py:144-175 written with torch ops (float32 on the CPU, the reference's own arithmetic order)."""
    from neuralangelo_b200 import glue
    g = torch.Generator().manual_seed(taps)
    shape = (2, 19, 7, 1)
    s = [torch.randn(shape, generator=g) * 0.3 for _ in range(taps)]
    sdf = torch.randn(shape, generator=g) * 0.3
    eps = (1.0 / 2049) / (np.sqrt(3) if taps == 4 else 1.0)
    sl = [t.clone().requires_grad_(True) for t in s]
    sd = sdf.clone().requires_grad_(True)
    if taps == 6:
        gradient = torch.cat([(sl[0] - sl[1]) / (2 * eps), (sl[2] - sl[3]) / (2 * eps), (sl[4] - sl[5]) / (2 * eps)], dim=-1)
        hessian = torch.cat([(sl[0] + sl[1] - 2 * sd) / (eps ** 2), (sl[2] + sl[3] - 2 * sd) / (eps ** 2),
                             (sl[4] + sl[5] - 2 * sd) / (eps ** 2)], dim=-1) if training else None
    else:
        k = torch.tensor([[1., -1., -1.], [-1., -1., 1.], [-1., 1., -1.], [1., 1., 1.]])
        gradient = (k[0] * sl[0] + k[1] * sl[1] + k[2] * sl[2] + k[3] * sl[3]) / (4.0 * eps)
        hxx = ((sl[0] + sl[1] + sl[2] + sl[3]) / 2.0 - 2 * sd) / eps ** 2
        hessian = torch.cat([hxx, hxx, hxx], dim=-1) / 3.0 if training else None
    wg, wh = torch.randn(*shape[:-1], 3, generator=g), torch.randn(*shape[:-1], 3, generator=g) * 1e-4
    loss = (gradient * wg).sum() + ((hessian * wh).sum() if training else 0.0)
    loss.backward()
    d = cuda_device
    tk = torch.stack(s).to(d).requires_grad_(True)
    sk = sdf.to(d).requires_grad_(True)
    grad_k, hess_k = glue.tap_combine(tk, sk, taps, eps, training)
    # same fp32 operations in the same order: equal to rounding of the division
    _close(grad_k, gradient, 3e-7, "gradient")
    if training:
        _close(hess_k, hessian, 3e-7, "hessian")
    else:
        assert hess_k is None
    lk = (grad_k * wg.to(d)).sum() + ((hess_k * wh.to(d)).sum() if training else 0.0)
    lk.backward()
    _close(tk.grad, torch.stack([t.grad for t in sl]), 2e-6, "d_taps")
    if training:
        _close(sk.grad, sd.grad, 2e-6, "d_sdf")


@pytest.mark.parametrize("with_hess,with_outside", [(True, True), (True, False), (False, True)])
def test_sdf_losses_match_reference_expression(na_lib, cuda_device, with_hess, with_outside):
    """misc.py:74-89: eikonal_loss / curvature_loss incl. nan_to_num and the outside mask."""
    from neuralangelo_b200 import glue
    from neuralangelo_b200.model import curvature_loss, eikonal_loss
    g = torch.Generator().manual_seed(3)
    B, R, N = 2, 50, 9
    grad = torch.randn(B, R, N, 3, generator=g) * 2
    hess = torch.randn(B, R, N, 3, generator=g) * 100
    hess[0, 1, 2] = float("inf")          # nan_to_num paths
    grad[1, 3, 4] = float("nan")
    outside = (torch.rand(B, R, 1, generator=g) < 0.2) if with_outside else None
    gd, hd = grad.double().requires_grad_(True), hess.double().requires_grad_(True)
    e_ref = eikonal_loss(gd, outside)
    c_ref = curvature_loss(hd, outside) if with_hess else None
    (3.0 * e_ref + (0.5 * c_ref if with_hess else 0.0)).backward()
    d = cuda_device
    gk, hk = grad.to(d).requires_grad_(True), hess.to(d).requires_grad_(True)
    e, c = glue.sdf_losses(gk, hk if with_hess else None, None if outside is None else outside.to(d))
    assert abs(e.item() - e_ref.item()) <= 2e-6 * abs(e_ref.item())
    if with_hess:
        assert abs(c.item() - c_ref.item()) <= 2e-6 * abs(c_ref.item())
    else:
        assert c is None
    (3.0 * e + (0.5 * c if with_hess else 0.0)).backward()
    finite = torch.isfinite(gd.grad).all(dim=-1)       # torch propagates NaN into the gradient of the NaN row itself
    _close(gk.grad.cpu()[finite], gd.grad[finite], 2e-5, "d_gradients")
    assert torch.isfinite(gk.grad).all()
    if with_hess:
        fin_h = torch.isfinite(hd.grad).all(dim=-1)
        _close(hk.grad.cpu()[fin_h], hd.grad[fin_h], 2e-6, "d_hessians")
