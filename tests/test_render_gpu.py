"""GPU parity: csrc/render.cu through the C ABI (via the reference-shaped wrappers) vs oracle/render_ref.py
and vs the golden fixtures produced by the reference's own code.  Tolerances: fp32, same formulas with a
different association order of the scans -> rtol 2e-5 / atol 2e-6 forward, 2e-4 relative on gradients."""
import os

import numpy as np
import pytest
import torch

from oracle import render_ref as rr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G(golden_dir):
    z = np.load(os.path.join(golden_dir, "render_golden.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def _scene(R, N, NB, seed, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    ray = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
    dists = torch.sort(torch.rand(R, N, generator=g) * 2 + 1, dim=1).values
    sdf = (2.0 - dists) * 0.4 + 0.02 * torch.randn(R, N, generator=g)
    grad = -ray[:, None, :] * (0.8 + 0.4 * torch.rand(R, N, 1, generator=g)) + 0.3 * torch.randn(R, N, 3, generator=g)
    far = dists[:, -1] + 0.05 + 0.1 * torch.rand(R, generator=g)
    rgb = torch.rand(R, N, 3, generator=g)
    dens = torch.rand(R, NB, generator=g) * 3
    dbg = far[:, None] / (1.0 - (torch.arange(NB) + torch.rand(R, NB, generator=g)) / NB + 1e-5)
    rgbb = torch.rand(R, NB, 3, generator=g)
    tgt = torch.rand(R, 3, generator=g)
    return [t.to(dtype) for t in (ray, dists, sdf, grad, far, rgb, dens, dbg, rgbb, tgt)]


def test_golden_reference_functions(G, na_lib, cuda_device):
    """The reference's own outputs (tests/golden) through the reference-shaped API."""
    from neuralangelo_b200 import render, nerf_util
    d = cuda_device
    w = render.alpha_compositing_weights(G["aw_alphas"].to(d))
    torch.testing.assert_close(w.cpu(), G["aw_weights"], rtol=2e-5, atol=1e-7)
    torch.testing.assert_close(render.composite(G["comp_q"].to(d), w).cpu(), G["comp_out"], rtol=2e-5, atol=2e-6)
    a = render.volume_rendering_alphas_dist(G["da_dens"].to(d), G["da_dists"].to(d), G["da_far"].to(d))
    torch.testing.assert_close(a.cpu(), G["da_alphas_far"], rtol=1e-5, atol=1e-6)
    a = render.volume_rendering_alphas_dist(G["da_dens"].to(d), G["da_dists"].to(d), None)
    torch.testing.assert_close(a.cpu(), G["da_alphas_nofar"], rtol=1e-5, atol=1e-6)
    out = nerf_util.sample_dists_from_pdf(G["pdf_bins"].to(d), G["pdf_w"].to(d), 16)
    torch.testing.assert_close(out.cpu(), G["pdf_out"], rtol=2e-5, atol=2e-5)
    s_var = torch.tensor(3.0, device=d)
    for tag, prog in (("p0", 0.0), ("p05", 0.05), ("p1", 1.0)):
        a = render.neus_alphas(G["na_ray"].to(d), G["na_sdf"].to(d), G["na_grad"].to(d), G["na_dists"].to(d), s_var,
                               dist_far=G["na_far"].to(d), progress=prog, anneal_end=0.1)
        torch.testing.assert_close(a.cpu(), G[f"na_alphas_{tag}"], rtol=2e-5, atol=2e-6)
    a = render.neus_alphas(G["na_ray"].to(d), G["na_sdf"].to(d), G["na_grad"].to(d), G["na_dists"].to(d), s_var)
    torch.testing.assert_close(a.cpu(), G["na_alphas_nofar"], rtol=2e-5, atol=2e-6)
    for h in (0, 2):
        f = nerf_util.sample_dists_hierarchical(G["na_dists"].to(d), G["na_sdf"].to(d), 64 * 2 ** h, 16)
        torch.testing.assert_close(f.cpu(), G[f"up_fine_h{h}"], rtol=2e-5, atol=2e-5)


def test_golden_backward_through_pieces_and_fused(G, na_lib, cuda_device):
    """d||rgb - tgt||^2 / d(sdf, grad, s_var) vs the reference's autograd -- both via the separate
    functions (alphas -> weights -> composite) and via the fused kernel."""
    from neuralangelo_b200 import render
    d = cuda_device
    for fused in (False, True):
        sdf = G["na_sdf"].to(d).requires_grad_(True)
        grad = G["na_grad"].to(d).requires_grad_(True)
        s_var = torch.tensor(3.0, device=d, requires_grad=True)
        args = (G["na_ray"].to(d), sdf, grad, G["na_dists"].to(d), s_var)
        if fused:
            rgb, _, _ = render.neus_render(*args, G["bw_rgbs"].to(d), dist_far=G["na_far"].to(d), progress=0.05)
        else:
            a = render.neus_alphas(*args, dist_far=G["na_far"].to(d), progress=0.05)
            rgb = render.composite(G["bw_rgbs"].to(d), render.alpha_compositing_weights(a))
        torch.testing.assert_close(rgb.detach().cpu(), G["bw_rgb"], rtol=2e-5, atol=2e-6)
        ((rgb - G["bw_tgt"].to(d)) ** 2).sum().backward()
        torch.testing.assert_close(sdf.grad.cpu(), G["bw_d_sdf"], rtol=2e-4, atol=2e-5)
        torch.testing.assert_close(grad.grad.cpu(), G["bw_d_grad"], rtol=2e-4, atol=2e-5)
        torch.testing.assert_close(s_var.grad.cpu(), G["bw_d_svar"], rtol=2e-4, atol=2e-5)


# >= 4096 rays with sample counts that are multiples of 4 take the shared-memory-staged kernel (bulk copies), the
# rest the direct one
@pytest.mark.parametrize("R,N,NB", [(257, 128, 32), (64, 37, 0), (33, 300, 17), (5, 1, 1), (4133, 128, 32), (4100, 64, 0),
                                    (4101, 36, 12), (4097, 200, 56)])
def test_fused_render_vs_oracle(na_lib, cuda_device, R, N, NB):
    from neuralangelo_b200 import render
    ray, dists, sdf, grad, far, rgb, dens, dbg, rgbb, tgt = _scene(R, N, max(NB, 1), seed=R)
    s = torch.tensor(2.5)
    leaf = lambda t: t.clone().requires_grad_(True)
    # oracle (CPU, float32 forward; float64 for gradients)
    o_in = [leaf(t.double()) for t in (sdf, grad, rgb, dens, rgbb)]
    so = s.double().requires_grad_(True)
    bgargs = dict(dens_bg=o_in[3], dists_bg=dbg.double(), rgb_bg=o_in[4]) if NB else {}
    o_rgb, o_w, o_op = rr.neus_render(o_in[0], o_in[1], dists.double(), ray.double(), far.double(), so, 0.3, o_in[2], **bgargs)
    ((o_rgb - tgt.double()) ** 2).sum().backward()
    # kernels
    d = cuda_device
    k_in = [leaf(t.to(d)) for t in (sdf, grad, rgb, dens, rgbb)]
    sk = s.to(d).requires_grad_(True)
    B = lambda t: t[None]  # add the batch dim the reference API carries
    kw = dict(densities_bg=B(k_in[3]), dists_bg=B(dbg.to(d))[..., None], rgbs_bg=B(k_in[4])) if NB else {}
    k_rgb, k_w, k_op = render.neus_render(B(ray.to(d)), B(k_in[0])[..., None], B(k_in[1]), B(dists.to(d))[..., None], sk,
                                          B(k_in[2]), dist_far=B(far.to(d))[..., None, None], progress=0.03, anneal_end=0.1,
                                          **kw)
    torch.testing.assert_close(k_rgb[0].cpu().double(), o_rgb.detach(), rtol=2e-5, atol=2e-6)
    torch.testing.assert_close(k_w[0, ..., 0].cpu().double(), o_w.detach(), rtol=2e-5, atol=2e-7)
    torch.testing.assert_close(k_op[0, ..., 0].cpu().double(), o_op.detach(), rtol=2e-5, atol=2e-6)
    ((k_rgb[0] - tgt.to(d)) ** 2).sum().backward()
    names = ["sdf", "grad", "rgb_obj"] + (["dens_bg", "rgb_bg"] if NB else [])
    for name, ko, oo in zip(names, k_in, o_in):
        scale = oo.grad.abs().max().item() + 1e-12
        err = (ko.grad.cpu().double() - oo.grad).abs().max().item()
        assert err <= 2e-4 * scale + 1e-7, (name, err, scale)
    assert abs(sk.grad.item() - so.grad.item()) <= 2e-4 * abs(so.grad.item()) + 1e-6


@pytest.mark.parametrize("want_weights,R", [(True, 97), (False, 97), (True, 4099)])
def test_white_background_gradient_through_opacity(na_lib, cuda_device, want_weights, R):
    """model.py:138-140 with background.white: rgb + (1 - opacity_all) back-propagates through the opacity
    (= sum of the compositing weights).  Oracle: float64 autograd of oracle/render_ref.py."""
    from neuralangelo_b200 import render
    N, NB = 64, 16
    ray, dists, sdf, grad, far, rgb, dens, dbg, rgbb, tgt = _scene(R, N, NB, seed=11)
    s = torch.tensor(2.5)
    leaf = lambda t: t.clone().requires_grad_(True)
    o_in = [leaf(t.double()) for t in (sdf, grad, rgb, dens, rgbb)]
    so = s.double().requires_grad_(True)
    o_rgb, o_w, o_op = rr.neus_render(o_in[0], o_in[1], dists.double(), ray.double(), far.double(), so, 0.3, o_in[2],
                                      dens_bg=o_in[3], dists_bg=dbg.double(), rgb_bg=o_in[4])
    o_white = o_rgb + (1 - o_op[..., None])
    ((o_white - tgt.double()) ** 2).sum().backward()
    d = cuda_device
    k_in = [leaf(t.to(d)) for t in (sdf, grad, rgb, dens, rgbb)]
    sk = s.to(d).requires_grad_(True)
    B = lambda t: t[None]
    k_rgb, k_w, k_op = render.neus_render(B(ray.to(d)), B(k_in[0])[..., None], B(k_in[1]), B(dists.to(d))[..., None], sk,
                                          B(k_in[2]), dist_far=B(far.to(d))[..., None, None], progress=0.03, anneal_end=0.1,
                                          densities_bg=B(k_in[3]), dists_bg=B(dbg.to(d))[..., None], rgbs_bg=B(k_in[4]),
                                          want_weights=want_weights)
    k_white = k_rgb + (1 - k_op)
    torch.testing.assert_close(k_white[0].cpu().double(), o_white.detach(), rtol=2e-5, atol=2e-6)
    ((k_white[0] - tgt.to(d)) ** 2).sum().backward()
    for name, ko, oo in zip(["sdf", "grad", "rgb_obj", "dens_bg", "rgb_bg"], k_in, o_in):
        scale = oo.grad.abs().max().item() + 1e-12
        err = (ko.grad.cpu().double() - oo.grad).abs().max().item()
        assert err <= 2e-4 * scale + 1e-7, (name, err, scale)
    assert abs(sk.grad.item() - so.grad.item()) <= 2e-4 * abs(so.grad.item()) + 1e-6


def test_alpha_weights_backward_with_saturated_alphas(na_lib, cuda_device):
    """alpha == 1 exactly (reachable through the clip at model.py:290): ATen's cumprod backward needs a
    special path here; the reverse affine scan must simply give the analytic gradient."""
    from neuralangelo_b200 import render
    a = torch.tensor([[[0.2, 1.0, 0.5, 0.0, 0.3]]], dtype=torch.float64, requires_grad=True)
    dw = torch.tensor([[[1.0, -2.0, 3.0, 0.5, 4.0]]], dtype=torch.float64)
    (rr.alpha_weights(a[0]) * dw[0]).sum().backward()
    ak = a.detach().float().to(cuda_device).requires_grad_(True)
    (render.alpha_compositing_weights(ak)[..., 0] * dw.float().to(cuda_device)).sum().backward()
    torch.testing.assert_close(ak.grad.cpu().double(), a.grad, rtol=1e-5, atol=1e-6)


def test_merge_sorted_and_sampling_loop(na_lib, cuda_device):
    from neuralangelo_b200 import nerf_util
    g = torch.Generator().manual_seed(4)
    R, N, NF = 130, 64, 16
    dists = torch.sort(torch.rand(1, R, N, 1, generator=g) * 2 + 1, dim=2).values
    sdfs = (2.0 - dists) * 0.4 + 0.02 * torch.randn(1, R, N, 1, generator=g)
    d = cuda_device
    dk, sk = dists.to(d), sdfs.to(d)
    do, so = dists[0, ..., 0], sdfs[0, ..., 0]
    for h in range(4):
        fine_k = nerf_util.sample_dists_hierarchical(dk, sk, 64 * 2 ** h, NF)
        fine_o = rr.neus_upsample(do, so, rr.midpoint_unif(NF), 64.0 * 2 ** h)
        torch.testing.assert_close(fine_k[0, ..., 0].cpu(), fine_o, rtol=2e-5, atol=2e-5)
        new_s = torch.randn(R, NF, generator=g) * 0.1
        # feed the ORACLE's fine samples to both sides so that rounding differences do not compound
        dk, sk = nerf_util.merge_sorted(dk, sk, fine_o[None, ..., None].to(d), new_s[None, ..., None].to(d))
        do, so = rr.merge_sorted(do, so, fine_o, new_s)
        assert torch.equal(dk[0, ..., 0].cpu(), do)
        assert torch.equal(sk[0, ..., 0].cpu(), so)
    assert dk.shape == (1, R, N + 4 * NF, 1)
    d2, none = nerf_util.merge_sorted(dists.to(d), None, fine_k)
    assert none is None and torch.all(d2[0, :, 1:, 0] >= d2[0, :, :-1, 0])


def test_large_ray_count_properties(na_lib, cuda_device):
    """2^18 rays x (128+32) samples (microbenchmark shape): weights sum to <= 1, equal the unfused path."""
    from neuralangelo_b200 import render
    d = cuda_device
    R, N, NB = 1 << 18, 128, 32
    g = torch.Generator(device=d).manual_seed(0)
    ray = torch.nn.functional.normalize(torch.randn(1, R, 3, device=d, generator=g), dim=-1)
    dists = torch.sort(torch.rand(1, R, N, 1, device=d, generator=g) * 2 + 1, dim=2).values
    sdf = (2.0 - dists) * 0.4
    grad = -ray[:, :, None, :].expand(1, R, N, 3).contiguous()
    rgb = torch.rand(1, R, N, 3, device=d, generator=g)
    dens = torch.rand(1, R, NB, device=d, generator=g)
    dbg = torch.sort(torch.rand(1, R, NB, 1, device=d, generator=g) * 10 + 3.2, dim=2).values
    rgbb = torch.rand(1, R, NB, 3, device=d, generator=g)
    s_var = torch.tensor(3.0, device=d)
    far = dists[:, :, -1:, :] + 0.1
    out, w, op = render.neus_render(ray, sdf, grad, dists, s_var, rgb, dist_far=far, densities_bg=dens, dists_bg=dbg,
                                    rgbs_bg=rgbb)
    a = torch.cat([render.neus_alphas(ray, sdf, grad, dists, s_var, dist_far=far),
                   render.volume_rendering_alphas_dist(dens, dbg)], dim=2)
    w2 = render.alpha_compositing_weights(a)
    out2 = render.composite(torch.cat([rgb, rgbb], dim=2), w2)
    assert torch.equal(w, w2) or (w - w2).abs().max().item() < 1e-6
    assert (out - out2).abs().max().item() < 1e-5
    assert op.max().item() <= 1.0 + 1e-5 and op.min().item() >= 0.0
