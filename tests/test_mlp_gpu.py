"""GPU parity of the tcgen05 fused SDF MLP (csrc/mlp_tcgen05.cu) against a plain PyTorch reference of the
same op (float64 on the CPU, and fp32 for scale).  Tolerance: the split-fp16 scheme is fp32-grade --
|err| <= 2e-6 * scale on the SDF (BASELINE.json asks <= 1e-3 relative; finite differences need far better)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mlp(k0, n, seed, beta=100.0):
    from neuralangelo_b200.mlp import MLPforNeuralSDF, get_activation
    torch.manual_seed(seed)
    mlp = MLPforNeuralSDF([k0, n, n], activ=get_activation("softplus", beta=beta), use_weightnorm=True,
                          geometric_init=True, out_bias=0.5)
    with torch.no_grad():  # make every weight matter (geometric init zeroes the encoding columns)
        for lin in mlp.linears:
            lin.weight_v.add_(torch.randn_like(lin.weight_v) * 0.05)
            lin.bias.add_(torch.randn_like(lin.bias) * 0.1)
    return mlp


@pytest.mark.parametrize("k0,n,m", [(131, 256, 1000), (131, 256, 128 * 148 * 2 + 77), (19, 64, 300), (67, 128, 129)])
def test_fused_sdf_mlp_forward(na_lib, cuda_device, k0, n, m):
    from neuralangelo_b200.fused_mlp import PackedSdfMlp
    mlp = _mlp(k0, n, seed=k0 + n)
    g = torch.Generator().manual_seed(m)
    x = torch.randn(m, k0, generator=g) * torch.tensor([1.0] * 3 + [0.3] * (k0 - 3))
    x[0, :] = 0.0
    x[1, 3:] = 1e-4          # tcnn-init-sized features (fp16 subnormal range for the high piece)
    ref_sdf, ref_feat = mlp.double()(x.double())
    mlp = mlp.float().to(cuda_device)
    packed = PackedSdfMlp(mlp, beta=100.0)
    buf = torch.zeros(m, k0 + 1, device=cuda_device)      # row stride k0+1 with a 1-float offset, like encode_rows
    buf[:, 1:] = x.to(cuda_device)
    sdf, feat, hidden = packed.forward(buf[:, 1:], with_sdf=True, with_feat=True, with_hidden=True)
    s_scale = ref_sdf.abs().max().item()
    assert (sdf.cpu().double() - ref_sdf).abs().max().item() <= 2e-6 * max(s_scale, 1.0)
    assert (feat.cpu().double() - ref_feat).abs().max().item() <= 2e-6 * max(ref_feat.abs().max().item(), 1.0)
    # sdf-only path (one layer) gives the same SDF
    sdf1, none = packed.forward(buf[:, 1:], with_sdf=True, with_feat=False)
    assert none is None and (sdf1 - sdf).abs().max().item() <= 1e-6 * max(s_scale, 1.0)   # head summed in another order
    # fp32 eager result of the module itself for scale: the fused kernel must be at least as accurate
    with torch.no_grad():
        e_sdf, e_feat = mlp(x.to(cuda_device))
    assert (sdf - e_sdf).abs().max().item() <= 5e-6 * max(s_scale, 1.0)


@pytest.mark.parametrize("k0,n,m", [(131, 256, 1), (131, 256, 127), (131, 256, 130), (131, 256, 128 * 148 * 3 + 5),
                                    (67, 128, 128 * 150 + 2), (35, 256, 1001)])
def test_pipelined_sdf_only_forward(na_lib, cuda_device, k0, n, m):
    """SDF-only evaluation takes the pipelined kernel (resident weights, two accumulator units, register-prefetched
    input tiles): a single short tile, ragged last tiles, several tiles per CTA, packed and strided rows.
    Reference: float64 module on the CPU; and the staged two-layer kernel, whose first layer runs the same sequence
    of MMAs per element (hidden activations equal up to the log1p evaluation: the pipelined epilogue alternates
    lg2.approx with an FMA-pipe polynomial)."""
    from neuralangelo_b200.fused_mlp import PackedSdfMlp
    mlp = _mlp(k0, n, seed=k0 + n + 1)
    g = torch.Generator().manual_seed(m)
    x = torch.randn(m, k0, generator=g) * torch.tensor([1.0] * 3 + [0.3] * (k0 - 3))
    ref_sdf, _ = mlp.double()(x.double(), with_feat=False)
    mlp = mlp.float().to(cuda_device)
    packed = PackedSdfMlp(mlp, beta=100.0)
    xp = x.to(cuda_device).contiguous()
    guard = torch.full((m + 64, 1), 7.0, device=cuda_device)       # the kernel must not write past row m
    sdf, _, hidden = packed.forward(xp, with_sdf=True, with_feat=False, with_hidden=True)
    assert (sdf.cpu().double() - ref_sdf).abs().max().item() <= 2e-6 * max(ref_sdf.abs().max().item(), 1.0)
    buf = torch.zeros(m, k0 + 1, device=cuda_device)
    buf[:, 1:] = xp
    sdf_s, _, hidden_s = packed.forward(buf[:, 1:], with_sdf=True, with_feat=False, with_hidden=True)
    assert torch.equal(hidden, hidden_s) and torch.equal(sdf, sdf_s)
    sdf_2, _, hidden_2 = packed.forward(xp, with_sdf=True, with_feat=True, with_hidden=True)   # staged kernel
    assert (hidden - hidden_2).abs().max().item() <= 1e-7 * max(hidden_2.abs().max().item(), 1.0)   # log1p: polynomial vs lg2
    assert (sdf - sdf_2).abs().max().item() <= 1e-6 * max(ref_sdf.abs().max().item(), 1.0)  # head sum order differs
    assert torch.all(guard == 7.0)
    # without the hidden activations base.yaml's shape takes the warp-specialised kernel (na_sdf_head_ws: loader warps,
    # pre-scaled operands, softplus in the log2 domain); other shapes the pipelined kernel again.  A second launch on the
    # same stream reproduces the first bit for bit, packed and strided rows agree bit for bit.
    scale = max(ref_sdf.abs().max().item(), 1.0)
    sdf2, _ = packed.forward(xp, with_sdf=True, with_feat=False)
    sdf3, _ = packed.forward(buf[:, 1:], with_sdf=True, with_feat=False)
    sdf4, _ = packed.forward(xp, with_sdf=True, with_feat=False)
    assert torch.equal(sdf2, sdf3) and torch.equal(sdf2, sdf4)
    assert (sdf2.cpu().double() - ref_sdf).abs().max().item() <= 2e-6 * scale
    if packed.ws_supported():
        assert (sdf2 - sdf).abs().max().item() <= 1e-6 * scale
    else:
        assert torch.equal(sdf2, sdf)


@pytest.mark.parametrize("k0,n,m", [(131, 256, 128 * 148 + 300), (131, 256, 77), (67, 128, 1000)])
def test_sdf_head_from_chunk_planes(na_lib, cuda_device, k0, n, m):
    """Chunk-plane input ([level][row][8] planes + one xyz chunk, the hash-grid kernels' native layout) gives the
    row-layout result: same products, K summed in a different order (features first, xyz last)."""
    from neuralangelo_b200.fused_mlp import PackedSdfMlp
    mlp = _mlp(k0, n, seed=k0 + n + 2)
    g = torch.Generator().manual_seed(m)
    x = torch.randn(m, k0, generator=g) * torch.tensor([1.0] * 3 + [0.3] * (k0 - 3))
    ref_sdf, _ = mlp.double()(x.double(), with_feat=False)
    mlp = mlp.float().to(cuda_device)
    packed = PackedSdfMlp(mlp, beta=100.0)
    xd = x.to(cuda_device)
    n_chunks = (k0 + 7) // 8
    m_alloc = m + 13                                   # chunk stride larger than the row count
    planes = torch.full((n_chunks, m_alloc, 8), float("nan"), device=cuda_device)
    planes[:n_chunks - 1, :m] = xd[:, 3:].reshape(m, n_chunks - 1, 8).permute(1, 0, 2)
    planes[n_chunks - 1, :m, :3] = xd[:, :3]          # the rest of the xyz chunk stays NaN: the kernel must ignore it
    sdf, hidden = packed.forward_planes(planes, m, with_hidden=True)
    scale = max(ref_sdf.abs().max().item(), 1.0)
    assert (sdf.cpu().double() - ref_sdf).abs().max().item() <= 2e-6 * scale
    sdf_r, _, hidden_r = packed.forward(xd.contiguous(), with_sdf=True, with_feat=False, with_hidden=True)
    assert (sdf - sdf_r).abs().max().item() <= 2e-6 * scale
    assert (hidden - hidden_r).abs().max().item() <= 2e-6 * max(hidden_r.abs().max().item(), 1.0)
    # without the hidden activations: the warp-specialised kernel at base.yaml's shape (na_sdf_head_ws, planes source)
    sdf_w = packed.forward_planes(planes, m)
    assert sdf_w.shape == (m, 1) and (sdf_w.cpu().double() - ref_sdf).abs().max().item() <= 2e-6 * scale
    assert torch.equal(sdf_w, packed.forward_planes(planes, m))


def test_one_launch_operand_packing_matches_the_torch_packing(na_lib, cuda_device):
    """na_pack_split_f16 (csrc/pack.cu) against fused_mlp.pack_weight (torch ops) bit for bit: plain, column-rotated
    (chunk-plane order), transposed with zero rows, scaled with the bias as an extra column."""
    import math
    from neuralangelo_b200.fused_mlp import PackedSdfMlp, pack_weight
    mlp = _mlp(131, 256, seed=3).float().to(cuda_device)
    with torch.no_grad():
        mlp.linears[0].weight_v[5, 7] = 1e6          # exercises the +-65504 clamp
    packed = PackedSdfMlp(mlp, beta=100.0)
    w0 = mlp.linears[0].effective_weight().detach().float()
    w1 = mlp.linears[1].effective_weight().detach().float()

    def same(name, w, k_pad):
        hi, lo = pack_weight(w, k_pad)
        assert torch.equal(getattr(packed, name + "_hi"), hi) and torch.equal(getattr(packed, name + "_lo"), lo), name

    same("w0", w0, 144)
    same("w1", w1, 256)
    same("w0p", torch.cat([w0[:, 3:], w0[:, :3]], dim=1), 144)
    w0t = torch.zeros(144, 256, device=cuda_device)
    w0t[:131] = w0.t()
    same("w0t", w0t, 256)
    same("w1t", w1.t().contiguous(), 256)
    c = 100.0 * math.log2(math.e)
    ws = torch.zeros(256, 144, dtype=torch.float64, device=cuda_device)
    ws[:, :128], ws[:, 128:131], ws[:, 131] = w0[:, 3:].double() * c, w0[:, :3].double() * c, packed.b0.double() * c
    same("w0s", ws.float(), 144)
    assert torch.equal(packed.w_head_s, packed.w_head * (math.log(2.0) / 100.0))


def test_finite_difference_gradient_quality(na_lib, cuda_device):
    """The reason for the split-precision scheme: central differences at eps = 1/2048 through the fused MLP
    must reproduce the float64 directional derivative (a TF32 single pass is off by O(1) here)."""
    from neuralangelo_b200.fused_mlp import PackedSdfMlp
    mlp = _mlp(131, 256, seed=5)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(4096, 131, generator=g) * 0.3
    d = torch.zeros(131)
    d[0] = 1.0
    eps = 1.0 / 2048
    md = mlp.double()
    ref = (md(x.double() + eps * d.double())[0] - md(x.double() - eps * d.double())[0]) / (2 * eps)
    mlp = mlp.float().to(cuda_device)
    packed = PackedSdfMlp(mlp, beta=100.0)
    xp, xm = (x + eps * d).to(cuda_device).contiguous(), (x - eps * d).to(cuda_device).contiguous()
    fd = (packed.forward(xp, with_feat=False)[0] - packed.forward(xm, with_feat=False)[0]) / (2 * eps)
    err = (fd.cpu().double() - ref).abs().max().item()
    assert err <= 2e-3 * max(ref.abs().max().item(), 1.0), err


@pytest.mark.parametrize("with_feat,m", [(True, 700), (False, 1300), (True, 8192)])   # 8192: split-K weight gradients
def test_fused_sdf_mlp_backward(na_lib, cuda_device, with_feat, m):
    """Gradients w.r.t. the input rows and every parameter (through weight norm) vs float64 autograd.
    Upstream gradients span 1e-9 .. 1e-3 per row (what a mean-reduced loss produces) to exercise the row scaling."""
    from neuralangelo_b200.fused_mlp import fused_sdf_mlp
    k0, n = 131, 256
    mlp = _mlp(k0, n, seed=11)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(m, k0, generator=g) * 0.3
    mag = 10.0 ** (-9 + 6 * torch.rand(m, 1, generator=g))
    d_sdf = torch.randn(m, 1, generator=g) * mag
    d_feat = torch.randn(m, n, generator=g) * mag * 0.1
    d_sdf[5] = 0.0
    d_feat[5] = 0.0                                      # an all-zero row
    # float64 reference
    md = mlp.double()
    xd = x.double().requires_grad_(True)
    s, f = md(xd, with_sdf=True, with_feat=with_feat)
    loss = (s * d_sdf.double()).sum() + ((f * d_feat.double()).sum() if with_feat else 0.0)
    loss.backward()
    ref = {k: p.grad.clone() for k, p in md.named_parameters() if p.grad is not None}
    ref_dx = xd.grad.clone()
    # fused
    mlp = mlp.float().to(cuda_device)
    for p in mlp.parameters():
        p.grad = None
    buf = torch.zeros(m, k0 + 1, device=cuda_device)
    buf[:, 1:] = x.to(cuda_device)
    rows = buf[:, 1:].requires_grad_(True)
    s, f = fused_sdf_mlp(mlp, rows, 100.0, with_sdf=True, with_feat=with_feat)
    loss = (s * d_sdf.to(cuda_device)).sum() + ((f * d_feat.to(cuda_device)).sum() if with_feat else 0.0)
    loss.backward()
    gx = rows.grad if rows.grad is not None else buf.grad[:, 1:]
    # per-row relative error of dx (rows differ by 6 orders of magnitude)
    err = (gx.cpu().double() - ref_dx).abs().amax(dim=1)
    scale = ref_dx.abs().amax(dim=1) + 1e-30
    assert (err / scale).max().item() <= 1e-4, (err / scale).max().item()
    for k, p in mlp.named_parameters():
        if k not in ref:
            assert p.grad is None or p.grad.abs().max().item() == 0, k
            continue
        r = ref[k]
        rel = (p.grad.cpu().double() - r).norm().item() / max(r.norm().item(), 1e-30)
        assert rel <= 2e-3, (k, rel)       # TF32 weight-gradient GEMMs


def _rgb_mlp(seed=0):
    from neuralangelo_b200.mlp import MLPwithSkipConnection
    import torch.nn.functional as F
    torch.manual_seed(seed)
    return MLPwithSkipConnection([278, 256, 256, 256, 256, 3], activ=F.relu, use_weightnorm=True)


def test_fused_chain_colour_mlp_forward_and_backward(na_lib, cuda_device):
    """278 -> 256 x4 -> 3 (+ sigmoid): input wider than the 256-column A operand (two accumulating products)."""
    from neuralangelo_b200 import fused_chain
    mlp = _rgb_mlp()
    g = torch.Generator().manual_seed(1)
    m = 128 * 3 + 5
    x = torch.randn(m, 278, generator=g) * 0.5
    tgt = torch.rand(m, 3, generator=g)
    md = mlp.double()
    xd = x.double().requires_grad_(True)
    ref = md(xd).sigmoid()
    ((ref - tgt.double()) ** 2).sum().backward()
    ref_g = {k: p.grad.clone() for k, p in md.named_parameters()}
    mlp = mlp.float().to(cuda_device)
    for p in mlp.parameters():
        p.grad = None
    xk = x.to(cuda_device).requires_grad_(True)
    with torch.no_grad():
        out_ng = fused_chain.mlp_forward(mlp, xk.detach(), final_act="sigmoid")
    assert (out_ng.cpu().double() - ref.detach()).abs().max().item() <= 2e-6
    out = fused_chain.fused_mlp(mlp, xk, final_act="sigmoid")
    assert torch.equal(out.detach(), out_ng)
    ((out - tgt.to(cuda_device)) ** 2).sum().backward()
    assert (xk.grad.cpu().double() - xd.grad).abs().max().item() <= 1e-5 * xd.grad.abs().max().item() + 1e-9
    for k, p in mlp.named_parameters():
        rel = (p.grad.cpu().double() - ref_g[k]).norm().item() / max(ref_g[k].norm().item(), 1e-30)
        # input-gradient chain: split-fp16 products (fp32-grade, checked on dx above); weight gradients: fp16 operands
        # with fp32 accumulation (11 significant bits, the class of the reference's TF32 products), biases fp32 sums
        assert rel <= (1e-5 if k.endswith("bias") else 2e-3), (k, rel)


def test_fused_chain_background_nerf(na_lib, cuda_device):
    """The whole BackgroundNeRF (8-layer skip MLP, 257-wide last layer split into density + feature, chained
    view-dependent colour MLP) in one launch vs the module's own torch forward."""
    from neuralangelo_b200 import fused_chain
    from neuralangelo_b200.config import default_model_config
    from neuralangelo_b200.modules import BackgroundNeRF
    torch.manual_seed(3)
    cfg = default_model_config()
    bg = BackgroundNeRF(cfg.background, appear_embed=cfg.appear_embed)
    g = torch.Generator().manual_seed(2)
    m = 1000
    pts = torch.randn(m, 3, generator=g) * 3 + 1.5
    rays = torch.nn.functional.normalize(torch.randn(m, 3, generator=g), dim=-1)
    bgd = bg.double()
    ref_rgb, ref_dens = bgd(pts.double(), rays.double(), None)
    bg = bg.float().to(cuda_device)
    with torch.no_grad():
        enc = bg.encode(pts.to(cuda_device)).contiguous()
        venc = bg.encode_view(rays.to(cuda_device)).contiguous()
        rgb, dens = fused_chain.background_forward(bg, enc, venc)
    assert (rgb.cpu().double() - ref_rgb).abs().max().item() <= 5e-6
    assert (dens.cpu().double() - ref_dens).abs().max().item() <= 5e-6 * max(ref_dens.abs().max().item(), 1.0)


@pytest.mark.parametrize("rows,cols", [(256, 278), (257, 84), (3, 256), (1, 256), (64, 131)])
def test_fused_weight_norm_forward_and_backward(na_lib, cuda_device, rows, cols):
    """csrc/weight_norm.cu against the torch expression of torch.nn.utils.weight_norm (dim=0) in float64."""
    from neuralangelo_b200.mlp import _WeightNorm
    g0 = torch.Generator().manual_seed(rows * 1000 + cols)
    v = torch.randn(rows, cols, generator=g0) * 0.3
    g = torch.rand(rows, 1, generator=g0) + 0.5
    dw = torch.randn(rows, cols, generator=g0)
    vd, gd = v.double().requires_grad_(True), g.double().requires_grad_(True)
    wd = vd * (gd / vd.norm(dim=1, keepdim=True))
    wd.backward(dw.double())
    vk, gk = v.to(cuda_device).requires_grad_(True), g.to(cuda_device).requires_grad_(True)
    wk = _WeightNorm.apply(vk, gk)
    wk.backward(dw.to(cuda_device))
    torch.testing.assert_close(wk.detach().cpu().double(), wd.detach(), rtol=2e-6, atol=1e-7)
    torch.testing.assert_close(vk.grad.cpu().double(), vd.grad, rtol=2e-5, atol=2e-6)
    torch.testing.assert_close(gk.grad.cpu().double(), gd.grad, rtol=2e-5, atol=2e-6)
    assert gk.grad.shape == (rows, 1)


@pytest.mark.parametrize("m,n,k", [(8192, 256, 131), (70001, 256, 278), (4096, 257, 340), (1000, 3, 256), (333, 1, 256),
                                   (65536, 128, 16), (50, 64, 351)])
def test_linear_wgrad_kernel_vs_float64(na_lib, cuda_device, m, n, k):
    """na_linear_wgrad (tcgen05, fp16 operands, fp32 accumulation) against a float64 product.  Tolerance: the operands
    carry 11 significant bits (as TF32), errors are independent over the m terms of a sum: 2e-3 of the largest entry."""
    from neuralangelo_b200.mlp import linear_wgrad
    g = torch.Generator(device=cuda_device).manual_seed(m + n + k)
    # gradients as the renderer produces them: tiny, with a heavy tail
    dy = torch.randn(m, n, device=cuda_device, generator=g) * 1e-6 * torch.exp(2 * torch.randn(m, 1, device=cuda_device, generator=g))
    x = torch.relu(torch.randn(m, k, device=cuda_device, generator=g)) * 3.0
    dw, db = linear_wgrad(dy, x)
    ref_w = dy.double().t() @ x.double()
    ref_b = dy.double().sum(0)
    assert dw.shape == (n, k) and db.shape == (n,)
    assert (dw.double() - ref_w).abs().max().item() <= 2e-3 * ref_w.abs().max().item()
    assert (db.double() - ref_b).abs().max().item() <= 2e-3 * ref_b.abs().max().item() + 1e-12
    # strided inputs (a column slice of a wider buffer), as the skip-connection layers pass them
    wide = torch.zeros(m, k + 5, device=cuda_device)
    wide[:, 2:2 + k] = x
    dw2, db2 = linear_wgrad(dy, wide[:, 2:2 + k])
    torch.testing.assert_close(dw2, dw, rtol=0, atol=1e-6 * dw.abs().max().item())
