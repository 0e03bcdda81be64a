"""CPU: pins oracle/render_ref.py against (a) fixtures generated from the reference's own code
(tests/golden/render_golden.npz, made by oracle/make_golden.py) and (b) the known-answer table of
SURVEY.md section 4."""
import os

import numpy as np
import pytest
import torch

from oracle import render_ref as rr


@pytest.fixture(scope="module")
def G(golden_dir):
    z = np.load(os.path.join(golden_dir, "render_golden.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def flat(t, keep=1):
    """[B,R,...] -> [B*R,...] dropping a trailing singleton dim when present."""
    t = t.reshape(-1, *t.shape[2:])
    return t


def test_kat_alpha_weights_and_composite():
    w = rr.alpha_weights(torch.tensor([[0.1, 0.5, 0.2, 0.9]]))
    np.testing.assert_allclose(w[0].numpy(), [0.1, 0.45, 0.09, 0.324], rtol=1e-6)
    np.testing.assert_allclose(w.sum().item(), 0.964, rtol=1e-6)


def test_kat_density_alphas():
    d = torch.tensor([[1.0, 2.0, 0.5]])
    t = torch.tensor([[0.5, 1.0, 2.0]])
    np.testing.assert_allclose(rr.density_alphas(d, t, None)[0].numpy(), [0.3934693, 0.8646647, 1.0], rtol=1e-6)
    np.testing.assert_allclose(rr.density_alphas(d, t, torch.tensor([3.0]))[0].numpy(),
                               [0.3934693, 0.8646647, 0.3934693], rtol=1e-6)


def test_kat_sample_pdf():
    bins = torch.tensor([[0., 1., 2., 3., 4.]])
    w = torch.tensor([[1., 0., 3., 0.]])
    np.testing.assert_allclose(rr.sample_pdf(bins, w, rr.midpoint_unif(4))[0].numpy(),
                               [0.5, 2.1666667, 2.5, 2.8333333], rtol=1e-6)
    np.testing.assert_allclose(rr.sample_pdf(bins, w, rr.midpoint_unif(8))[0].numpy(),
                               [0.25, 0.75, 2.0833333, 2.25, 2.4166667, 2.5833333, 2.75, 2.9166667], rtol=1e-6)
    np.testing.assert_allclose(rr.sample_pdf(bins, torch.zeros(1, 4), rr.midpoint_unif(4))[0].numpy(), [4, 4, 4, 4])


def test_kat_neus_alphas():
    sdf = torch.tensor([[.3, .1, -.1, -.3]])
    grad = torch.tensor([0.6, 0., -0.5]).expand(1, 4, 3)
    dists = torch.tensor([[1, 1.2, 1.4, 1.6]])
    ray = torch.tensor([[0., 0., 1.]])
    far = torch.tensor([1.8])
    s = torch.tensor(3.0)
    for ratio, expect in ((0.0, [0.0102499, 0.3585021, 0.9233539, 0.9494578]),
                          (0.5, [0.0077237, 0.2941217, 0.8849208, 0.9170649]),
                          (1.0, [0.0056729, 0.2321243, 0.8252234, 0.8637335])):
        np.testing.assert_allclose(rr.neus_alphas(sdf, grad, dists, ray, far, s, ratio)[0].numpy(), expect, rtol=2e-5)
    assert abs(rr.neus_alphas(sdf, grad, dists, ray, None, s, 1.0)[0, -1].item() - 0.99999) < 1e-5


def test_kat_upsample(G):
    sdf = torch.tensor([[.3, .1, -.1, -.3]])
    dists = torch.tensor([[1, 1.2, 1.4, 1.6]])
    got = rr.neus_upsample(dists, sdf, rr.midpoint_unif(4), 64.0)
    np.testing.assert_allclose(got[0].numpy(), [1.2247505, 1.2749169, 1.3250834, 1.3752497], rtol=1e-6)
    np.testing.assert_allclose(got[0].numpy(), G["kat_up"].reshape(-1).numpy(), rtol=1e-6)


def test_golden_alpha_weights_composite(G):
    a = flat(G["aw_alphas"])
    w = rr.alpha_weights(a)
    torch.testing.assert_close(w, flat(G["aw_weights"])[..., 0], rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(rr.composite(flat(G["comp_q"]), w), flat(G["comp_out"]), rtol=1e-5, atol=1e-6)


def test_golden_density_alphas(G):
    dens, dists, far = flat(G["da_dens"]), flat(G["da_dists"])[..., 0], flat(G["da_far"]).reshape(-1)
    torch.testing.assert_close(rr.density_alphas(dens, dists, far), flat(G["da_alphas_far"]), rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(rr.density_alphas(dens, dists, None), flat(G["da_alphas_nofar"]), rtol=1e-6, atol=1e-7)


def test_golden_sample_pdf(G):
    bins, w = flat(G["pdf_bins"])[..., 0], flat(G["pdf_w"])
    got = rr.sample_pdf(bins, w, rr.midpoint_unif(16))
    torch.testing.assert_close(got, flat(G["pdf_out"])[..., 0], rtol=1e-5, atol=1e-5)


def _neus_inputs(G):
    return (flat(G["na_sdf"])[..., 0], flat(G["na_grad"]), flat(G["na_dists"])[..., 0], flat(G["na_ray"]),
            flat(G["na_far"]).reshape(-1))


def test_golden_neus_alphas(G):
    sdf, grad, dists, ray, far = _neus_inputs(G)
    s = torch.tensor(3.0)
    for tag, ratio in (("p0", 0.0), ("p05", 0.5), ("p1", 1.0)):  # progress / anneal_end(0.1), clipped at 1
        torch.testing.assert_close(rr.neus_alphas(sdf, grad, dists, ray, far, s, ratio), flat(G[f"na_alphas_{tag}"]),
                                   rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(rr.neus_alphas(sdf, grad, dists, ray, None, s, 1.0), flat(G["na_alphas_nofar"]),
                               rtol=1e-5, atol=1e-6)


def test_golden_render_backward(G):
    """Gradients of ||rgb - target||^2 through alphas -> weights -> composite, vs the reference's autograd."""
    sdf, grad, dists, ray, far = _neus_inputs(G)
    sdf, grad = sdf.clone().requires_grad_(True), grad.clone().requires_grad_(True)
    s = torch.tensor(3.0, requires_grad=True)
    rgb, _, _ = rr.neus_render(sdf, grad, dists, ray, far, s, 0.5, flat(G["bw_rgbs"]))
    torch.testing.assert_close(rgb, flat(G["bw_rgb"]), rtol=1e-5, atol=1e-6)
    ((rgb - flat(G["bw_tgt"])) ** 2).sum().backward()
    torch.testing.assert_close(sdf.grad, flat(G["bw_d_sdf"])[..., 0], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(grad.grad, flat(G["bw_d_grad"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(s.grad, G["bw_d_svar"], rtol=1e-4, atol=1e-5)


def test_golden_upsample(G):
    _, _, dists, _, _ = _neus_inputs(G)
    sdf = flat(G["na_sdf"])[..., 0]
    for h in (0, 2):
        got = rr.neus_upsample(dists, sdf, rr.midpoint_unif(16), 64.0 * 2 ** h)
        torch.testing.assert_close(got, flat(G[f"up_fine_h{h}"])[..., 0], rtol=1e-5, atol=1e-5)


def test_merge_sorted_matches_sort():
    g = torch.Generator().manual_seed(0)
    a = torch.sort(torch.rand(7, 20, generator=g), dim=1).values
    b = torch.sort(torch.rand(7, 6, generator=g), dim=1).values
    sa, sb = torch.randn(7, 20, generator=g), torch.randn(7, 6, generator=g)
    d, s = rr.merge_sorted(a, sa, b, sb)
    assert torch.all(d[:, 1:] >= d[:, :-1])
    lookup = {float(k): float(v) for k, v in zip(torch.cat([a, b], 1).flatten(), torch.cat([sa, sb], 1).flatten())}
    for k, v in zip(d.flatten(), s.flatten()):
        assert lookup[float(k)] == float(v)
