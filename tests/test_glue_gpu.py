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


def test_ray_setup_matches_the_reference_formulation(na_lib, cuda_device):
    """na_ray_setup (pixel index -> centre, ray, unit ray, unit-sphere bounds; one launch) against the torch formulation of
    camera.py:283-311 + model.py:117-121,207-213 evaluated in float64, for cameras whose rays hit and miss the sphere."""
    from neuralangelo_b200 import camera, nerf_util
    g = torch.Generator().manual_seed(3)
    B, R, H, W = 3, 4096, 600, 800
    poses = []
    for b in range(B):
        eye = torch.randn(3, generator=g, dtype=torch.float64)
        eye = eye / eye.norm() * (2.0 + b)
        fwd = -eye / eye.norm()
        right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0], dtype=torch.float64), fwd)
        right = right / right.norm()
        Rm = torch.stack([right, torch.linalg.cross(fwd, right), fwd])
        poses.append(torch.cat([Rm, (-Rm @ eye)[:, None]], dim=1))
    pose = torch.stack(poses)
    intr = torch.tensor([[700.0, 0.3, 400.0], [0.0, 710.0, 300.0], [0.0, 0.0, 1.0]], dtype=torch.float64).repeat(B, 1, 1)
    idx = torch.randint(0, H * W, (B, R), generator=g)
    xy = torch.stack([(idx % W).double() + 0.5, torch.div(idx, W, rounding_mode="floor").double() + 0.5], dim=-1)
    center64, ray64 = camera._rays_from_pixels(xy, pose, intr)
    unit64 = torch.nn.functional.normalize(ray64, dim=-1)
    near64, far64 = nerf_util.intersect_with_sphere(center64, unit64, radius=1.0)
    c, r, u, (near, far, outside) = camera.ray_setup(pose.float().to(cuda_device), intr.float().to(cuda_device), (H, W),
                                                     idx.to(cuda_device))
    assert c.shape == (B, R, 3) and near.shape == (B, R, 1) and outside.dtype == torch.bool
    assert (c.cpu().double() - center64).abs().max().item() <= 2e-6 * center64.abs().max().item()
    assert (r.cpu().double() - ray64).abs().max().item() <= 5e-6 * ray64.abs().max().item()   # world point minus centre
    assert (u.cpu().double() - unit64).abs().max().item() <= 5e-6
    disc2 = ((center64 * unit64).sum(-1, keepdim=True) ** 2 - ((center64 * center64).sum(-1, keepdim=True) - 1.0))
    hit, miss = disc2 > 1e-3, disc2 < -1e-3
    assert hit.any() and miss.any()
    assert bool((outside.cpu() == near64.isnan())[hit | miss].all())
    assert (near.cpu().double() - near64.clamp_min(0.0))[hit].abs().max().item() <= 2e-4
    assert (far.cpu().double() - far64)[hit].abs().max().item() <= 2e-4
    assert bool((near.cpu()[miss] == 1.0).all()) and bool((far.cpu()[miss] == torch.tensor(1.2).float()).all())
    # no backward: differentiable cameras must take the torch formulation
    with pytest.raises(NotImplementedError):
        camera.ray_setup(pose.float().to(cuda_device).requires_grad_(), intr.float().to(cuda_device), (H, W), idx.to(cuda_device))
