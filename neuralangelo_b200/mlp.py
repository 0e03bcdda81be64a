"""Weight-normed skip MLPs of the Neuralangelo path, with the reference's module/parameter names so that
its checkpoints load unchanged (`linears.{i}.{weight_g,weight_v,bias}`, `linear_sdf.{weight,bias}`).

Mirrors `projects/neuralangelo/utils/mlp.py:18-84` (MLPforNeuralSDF) and
`projects/nerf/utils/nerf_util.py:158-196` (MLPwithSkipConnection).  The weight-norm reparametrisation
W = g * v / ||v|| is evaluated once per forward of the whole MLP (the reference's hook recomputes it on
every `linear(...)` call, >= 7 times per step for the SDF MLP).
"""
import numpy as np
import torch
import torch.nn.functional as F


def get_activation(activ, **kwargs):
    """projects/neuralangelo/utils/misc.py:92-107 (the activations the path uses)."""
    if activ in ("relu", "relu_"):
        return F.relu
    if activ == "softplus":
        beta = float(kwargs.get("beta", 1.0))
        return lambda x: F.softplus(x, beta=beta)
    if activ == "identity":
        return lambda x: x
    if activ in ("sigmoid", "sigmoid_"):
        return torch.sigmoid
    if activ in ("silu", "silu_"):
        return F.silu
    if activ in ("abs", "abs_"):
        return torch.abs
    if activ in ("exp", "exp_"):
        return torch.exp
    raise KeyError(activ)


def tn_matmul(a, b):
    """a^T b for tall operands a [M,p], b [M,q] (M >> p, q: a weight gradient, i.e. a reduction over all rows).  As one
    GEMM the output has only a handful of tiles and the library runs it on a few SMs (measured on B200: 312 us for
    M = 262144, p = 256, q = 131); split along M into independent batched products it fills the GPU."""
    m = a.shape[0]
    for parts in (64, 32, 16, 8):
        if m % parts == 0 and m // parts >= 1024:
            return torch.bmm(a.reshape(parts, m // parts, a.shape[1]).transpose(1, 2),
                             b.reshape(parts, m // parts, b.shape[1])).sum(dim=0)
    return a.t() @ b


def linear_wgrad(dy, x):
    """(dW [N,K], db [N]) of y = x W^T + b for dy [M,N], x [M,K] on the tcgen05 weight-gradient kernel
    (na_linear_wgrad): one launch, fp16 operands (TF32-class mantissa) with dy pre-scaled by a power of two that each CTA
    takes from the largest magnitude in its own row slice, fp32 accumulation."""
    from . import _lib
    m, n = dy.shape
    k = x.shape[1]
    dy = dy if dy.stride(1) == 1 else dy.contiguous()
    x = x if x.stride(1) == 1 else x.contiguous()
    ld_w = (k + 1 + 3) // 4 * 4
    out = torch.zeros(n, ld_w, dtype=torch.float32, device=dy.device)
    # amax = NULL: every CTA derives its power-of-two scale from a sweep over its own slice of dy
    _lib.check(_lib.lib().na_linear_wgrad(_lib.ptr(dy), dy.stride(0), _lib.ptr(x), x.stride(0), m, n, k, None,
                                          _lib.ptr(out), ld_w, _lib.current_stream()), "na_linear_wgrad")
    return out[:, :k], out[:, k]


def wgrad_supported(dy, x):
    return dy.is_cuda and dy.dtype == torch.float32 and x.dtype == torch.float32 and x.shape[1] <= 352 and \
        dy.dim() == 2 and x.dim() == 2


def colsum(a):
    """a.sum(dim=0) of a tall contiguous [M,p] matrix (a bias gradient).  ATen's reduction over the slow dimension
    runs at a quarter of the HBM bandwidth for M >= 65536 (measured on B200: 168 us for 262144 x 256 fp32); summing 64
    row blocks first and the 64 partial rows afterwards takes 65 us (and rounds less)."""
    m = a.shape[0]
    if a.is_contiguous() and m % 64 == 0 and m >= 4096:
        return a.view(64, m // 64, a.shape[1]).sum(dim=1).sum(dim=0)
    return a.sum(dim=0)


class _LinearTallRows(torch.autograd.Function):
    """y = x W^T + b with the weight gradient evaluated by `tn_matmul` (the forward and dx are the library GEMMs
    F.linear would run; only dW, a [out,in] result reduced over >= 8192 rows, is re-shaped into a split-K batch)."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w)
        return torch.addmm(b, x, w.t())

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx = dy @ w if ctx.needs_input_grad[0] else None
        if (ctx.needs_input_grad[1] or ctx.needs_input_grad[2]) and wgrad_supported(dy, x):
            dw, db = linear_wgrad(dy, x)        # tcgen05 kernel: both gradients in one launch
            return dx, (dw if ctx.needs_input_grad[1] else None), (db if ctx.needs_input_grad[2] else None)
        dw = tn_matmul(dy, x) if ctx.needs_input_grad[1] else None
        db = colsum(dy) if ctx.needs_input_grad[2] else None
        return dx, dw, db


def linear(x, w, b):
    """F.linear; on CUDA with many rows and a graph to build, the variant whose weight gradient is split-K.  `x` may carry
    ZERO columns beyond the layer's input width (glue.rgb_input pads 278 -> 280 so that K is a multiple of four and the
    library picks its vectorised GEMMs: 30 us instead of 112 us forward, 36 instead of 126 us for dx at 65 536 rows): the
    weight is zero-padded to match, autograd slices its gradient back."""
    if x.shape[-1] > w.shape[-1]:
        w = F.pad(w, (0, x.shape[-1] - w.shape[-1]))
    if x.is_cuda and x.dim() == 2 and x.shape[0] >= 1024 and b is not None and torch.is_grad_enabled() and \
            (w.requires_grad or x.requires_grad):
        return _LinearTallRows.apply(x, w, b)
    return F.linear(x, w, b)


class _WeightNorm(torch.autograd.Function):
    """W = g * v / ||v|| per output row as ONE launch forward and one backward (csrc/weight_norm.cu) instead of the
    ~nine ATen kernels of the composite expression; used on CUDA only (the CPU path keeps the torch expression)."""

    @staticmethod
    def forward(ctx, v, g):
        from . import _lib
        vc, gc = v.detach().contiguous(), g.detach().reshape(-1).contiguous()
        rows, cols = vc.shape
        w = torch.empty_like(vc)
        inv = torch.empty(rows, dtype=torch.float32, device=vc.device)
        _lib.check(_lib.lib().na_weight_norm_fwd(_lib.ptr(vc), _lib.ptr(gc), rows, cols, _lib.ptr(w), _lib.ptr(inv),
                                                 _lib.current_stream()), "na_weight_norm_fwd")
        ctx.save_for_backward(vc, gc, inv)
        ctx.g_shape = g.shape
        return w

    @staticmethod
    def backward(ctx, dw):
        from . import _lib
        vc, gc, inv = ctx.saved_tensors
        rows, cols = vc.shape
        dwc = dw.contiguous()
        dv, dg = torch.empty_like(vc), torch.empty_like(gc)
        _lib.check(_lib.lib().na_weight_norm_bwd(_lib.ptr(vc), _lib.ptr(gc), _lib.ptr(inv), _lib.ptr(dwc), rows, cols,
                                                 _lib.ptr(dv), _lib.ptr(dg), _lib.current_stream()), "na_weight_norm_bwd")
        return dv, dg.view(ctx.g_shape)


class Linear(torch.nn.Module):
    """nn.Linear, optionally weight-normed with the legacy `weight_g` / `weight_v` parameter names of
    `torch.nn.utils.weight_norm` (dim=0: one norm per output row)."""

    def __init__(self, k_in, k_out, weight_norm=False):
        super().__init__()
        self.in_features, self.out_features, self.weight_norm = k_in, k_out, weight_norm
        w = torch.empty(k_out, k_in)
        torch.nn.init.kaiming_uniform_(w, a=np.sqrt(5))
        bound = 1 / np.sqrt(k_in)
        self.bias = torch.nn.Parameter(torch.empty(k_out).uniform_(-bound, bound))
        if weight_norm:
            self.weight_g = torch.nn.Parameter(w.norm(dim=1, keepdim=True))
            self.weight_v = torch.nn.Parameter(w)
        else:
            self.weight = torch.nn.Parameter(w)

    def set_weight(self, w):
        """Assign the effective weight (used by the initialisers, which run before weight-norm in the reference)."""
        with torch.no_grad():
            if self.weight_norm:
                self.weight_v.copy_(w)
                self.weight_g.copy_(w.norm(dim=1, keepdim=True))
            else:
                self.weight.copy_(w)

    def effective_weight(self):
        if not self.weight_norm:
            return self.weight
        if self.weight_v.is_cuda and self.weight_v.dtype == torch.float32:
            return _WeightNorm.apply(self.weight_v, self.weight_g)
        return self.weight_v * (self.weight_g / self.weight_v.norm(dim=1, keepdim=True))

    def forward(self, x):
        if x.dim() > 2 and x.is_cuda:
            lead = x.shape[:-1]
            return linear(x.reshape(-1, x.shape[-1]), self.effective_weight(), self.bias).view(*lead, -1)
        return linear(x, self.effective_weight(), self.bias)


class MLPforNeuralSDF(torch.nn.Module):
    """mlp.py:18-84.  forward(input, with_sdf, with_feat) -> [sdf | None, feat | None]; the SDF head reads the
    INPUT of the last feature layer (mlp.py:66-67)."""

    def __init__(self, layer_dims, skip_connection=(), activ=None, use_layernorm=False, use_weightnorm=False,
                 geometric_init=False, out_bias=0., invert=False):
        super().__init__()
        if use_layernorm:
            raise NotImplementedError("layer norm is not used on the Neuralangelo path (base.yaml)")
        self.skip_connection = list(skip_connection)
        self.linears = torch.nn.ModuleList()
        pairs = list(zip(layer_dims[:-1], layer_dims[1:]))
        for li, (k_in, k_out) in enumerate(pairs):
            if li in self.skip_connection:
                k_in += layer_dims[0]
            lin = Linear(k_in, k_out, weight_norm=use_weightnorm)
            if geometric_init:  # mlp.py:71-77
                w = torch.empty(k_out, k_in).normal_(0.0, np.sqrt(2 / k_out))
                if li == 0:
                    w[:, 3:] = 0.0
                if li in self.skip_connection:
                    w[:, -layer_dims[0]:] = 0.0
                lin.set_weight(w)
                torch.nn.init.constant_(lin.bias, 0.0)
            if li == len(pairs) - 1:
                torch.nn.init.constant_(lin.bias, 0.0)
            self.linears.append(lin)
        self.linear_sdf = Linear(k_in, 1)
        if geometric_init:  # mlp.py:79-84
            with torch.no_grad():
                self.linear_sdf.weight.normal_(mean=np.sqrt(np.pi / k_in), std=0.0001)
                self.linear_sdf.bias.fill_(-out_bias)
                if invert:
                    self.linear_sdf.weight.mul_(-1)
                    self.linear_sdf.bias.mul_(-1)
        self.activ = activ or F.relu

    def forward(self, input, with_sdf=True, with_feat=True):
        feat = input
        n = len(self.linears)
        for li, linear in enumerate(self.linears):
            if li in self.skip_connection:
                feat = torch.cat([feat, input], dim=-1)
            last = li == n - 1
            if last:
                sdf = self.linear_sdf(feat) if with_sdf else None
                out_feat = self.activ(linear(feat)) if with_feat else None
                return [sdf, out_feat]
            feat = self.activ(linear(feat))


class MLPwithSkipConnection(torch.nn.Module):
    """nerf_util.py:158-196."""

    def __init__(self, layer_dims, skip_connection=(), activ=None, use_layernorm=False, use_weightnorm=False):
        super().__init__()
        if use_layernorm:
            raise NotImplementedError("layer norm is not used on the Neuralangelo path (base.yaml)")
        self.skip_connection = list(skip_connection)
        self.linears = torch.nn.ModuleList()
        pairs = list(zip(layer_dims[:-1], layer_dims[1:]))
        for li, (k_in, k_out) in enumerate(pairs):
            if li in self.skip_connection:
                k_in += layer_dims[0]
            lin = Linear(k_in, k_out, weight_norm=use_weightnorm)
            if li == len(pairs) - 1:
                torch.nn.init.constant_(lin.bias, 0.0)
            self.linears.append(lin)
        self.activ = activ or F.relu

    def forward(self, input):
        feat = input
        for li, linear in enumerate(self.linears):
            if li in self.skip_connection:
                feat = torch.cat([feat, input], dim=-1)
            feat = linear(feat)
            if li != len(self.linears) - 1:
                feat = self.activ(feat)
        return feat
