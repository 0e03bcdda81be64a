"""Host side of the generic tcgen05 MLP chain (csrc/mlp_chain.cu): builds the stage programs for
`MLPwithSkipConnection` (colour network) and for the whole `BackgroundNeRF` (feature MLP -> density + view-dependent
colour MLP in ONE launch), and wraps them for autograd.

Forward runs fused on the tensor cores with the split-fp16 arithmetic; for training the forward also saves each
layer's activations and the backward is evaluated with library GEMMs on those (TF32 like the reference,
imaginaire/trainers/base.py:172-175) -- the colour / background networks do not feed finite differences, so they
do not need the fp32-grade input-gradient kernel the SDF network has.
"""
import ctypes

import torch

from . import _lib
from ._lib import MlpStage, ptr
from .fused_mlp import pack_weight, pad16

_L = _lib.lib
ACT = {"none": 0, "relu": 1, "softplus": 2, "sigmoid": 3}
MAX_K = 256


class _Program:
    """Accumulates stages; keeps every tensor whose address is baked into the program alive."""

    def __init__(self):
        self.stages, self.keep = [], []
        self._cur = MlpStage()

    def epilogue(self, bias, act, n_valid, to_a, out=None, save=None, act_param=0.0):
        s = self._cur
        b = bias.detach().float().contiguous()
        self.keep.append(b)
        s.has_epi, s.bias, s.act, s.act_param, s.n_valid, s.epi_to_a = 1, b.data_ptr(), ACT[act], float(act_param), n_valid, int(to_a)
        if out is not None:
            assert out.stride(-1) == 1
            s.out, s.out_ld = out.data_ptr(), out.stride(0)
            self.keep.append(out)
        if save is not None:
            s.save, s.save_ld = save.data_ptr(), save.stride(0)
            self.keep.append(save)

    def load(self, x, col, ncols):
        assert ncols <= MAX_K and x.stride(-1) == 1 and x.dtype == torch.float32
        s = self._cur
        s.inp, s.in_ld, s.in_col, s.in_cols = x.data_ptr(), x.stride(0), col, ncols
        self.keep.append(x)

    def gemm(self, w_seg, accumulate):
        """w_seg: effective weight block [n_out, k] multiplying the columns currently staged in A."""
        n_out, k = w_seg.shape
        n_pad, k_pad = pad16(n_out), pad16(k)
        wp = torch.zeros(n_pad, k_pad, dtype=torch.float32, device=w_seg.device)
        wp[:n_out, :k] = w_seg.detach().float()
        hi, lo = pack_weight(wp, k_pad)
        self.keep += [hi, lo]
        s = self._cur
        s.has_gemm, s.w_hi, s.w_lo, s.k_chunks, s.n_out, s.accumulate = 1, hi.data_ptr(), lo.data_ptr(), k_pad // 8, n_pad, int(accumulate)
        self._flush()

    def end(self):
        self._flush()

    def _flush(self):
        self.stages.append(self._cur)
        self._cur = MlpStage()

    def run(self, m):
        arr = (MlpStage * len(self.stages))(*self.stages)
        _lib.check(_L().na_mlp_chain_fwd(arr, len(self.stages), m, _lib.current_stream()), "na_mlp_chain_fwd")


def _input_segments(k0):
    return [(c, min(MAX_K, k0 - c)) for c in range(0, k0, MAX_K)]


def _append_mlp(prog, mlp, x, hidden_act, saves=None, first_epilogue=None):
    """Appends every product of an MLPwithSkipConnection to `prog`.  The network input is either the global tensor
    `x` (staged in <= 256-column segments) or, when `first_epilogue` is given, the activation a preceding epilogue
    leaves in the A operand followed by the columns of `x` (a [A-resident | x] concatenation).  Returns
    (bias, n_out) of the last layer, whose epilogue the caller still has to emit."""
    layers = list(mlp.linears)
    k0 = x.shape[1]
    for li, lin in enumerate(layers):
        w = lin.effective_weight()
        if li == 0:
            if first_epilogue is None:
                for c, n in _input_segments(k0):
                    prog.load(x, c, n)
                    prog.gemm(w[:, c:c + n], accumulate=c > 0)
            else:
                resident = first_epilogue()            # emits the epilogue that fills A; returns its width
                prog.gemm(w[:, :resident], accumulate=False)
                for c, n in _input_segments(k0):
                    prog.load(x, c, n)
                    prog.gemm(w[:, resident + c:resident + c + n], accumulate=True)
            continue
        prev = layers[li - 1]
        n_prev = prev.out_features
        assert n_prev <= MAX_K, "hidden width > 256 is outside the fused kernel's envelope"
        prog.epilogue(prev.bias, hidden_act, n_prev, to_a=True, save=None if saves is None else saves[li - 1])
        prog.gemm(w[:, :n_prev], accumulate=False)
        if li in mlp.skip_connection:      # feat = cat([feat, input]) before this layer (nerf_util.py:189-190)
            assert first_epilogue is None
            for c, n in _input_segments(k0):
                prog.load(x, c, n)
                prog.gemm(w[:, n_prev + c:n_prev + c + n], accumulate=True)
    return layers[-1].bias, layers[-1].out_features


def _act_name(fn):
    import torch.nn.functional as F
    if fn in (F.relu, F.relu_, torch.relu):
        return "relu"
    raise NotImplementedError("the fused chain covers relu hidden activations (base.yaml: rgb relu_, background relu)")


def _supported(mlp):
    try:
        _act_name(mlp.activ)
    except NotImplementedError:
        return False
    return all(l.out_features <= MAX_K for l in mlp.linears)


def mlp_forward(mlp, x, final_act="none", saves=None):
    """Fused forward of an MLPwithSkipConnection on rows x [M,K0] (fp32, unit inner stride) -> [M, n_out]."""
    _lib.require_cuda(x)
    m = x.shape[0]
    prog = _Program()
    bias, n_out = _append_mlp(prog, mlp, x, _act_name(mlp.activ), saves=saves)
    out = torch.empty(m, n_out, dtype=torch.float32, device=x.device)
    prog.epilogue(bias, final_act, n_out, to_a=False, out=out)
    prog.end()
    prog.run(m)
    return out


def background_forward(bg, points_enc, view_enc):
    """Whole BackgroundNeRF network (modules.py:270-285) in one launch: mlp_feat -> (softplus density | relu
    feature) -> mlp_rgb([feature, view_enc]) -> sigmoid.  Returns (rgb [M,3], density [M])."""
    _lib.require_cuda(points_enc, view_enc)
    m = points_enc.shape[0]
    feat_mlp, rgb_mlp = bg.mlp_feat, bg.mlp_rgb
    hidden = _act_name(feat_mlp.activ)
    prog = _Program()
    _append_mlp(prog, _Truncated(feat_mlp), points_enc, hidden)     # all but the last (257-wide) layer
    last = feat_mlp.linears[-1]
    w_last = last.effective_weight()
    prev = feat_mlp.linears[-2]
    # last layer of mlp_feat as two products of the same A operand: column 0 (density), then columns 1.. (feature)
    prog.epilogue(prev.bias, hidden, prev.out_features, to_a=True)
    prog.gemm(w_last[:1], accumulate=False)
    density = torch.empty(m, 1, dtype=torch.float32, device=points_enc.device)
    prog.epilogue(last.bias[:1], "softplus", 1, to_a=False, out=density, act_param=1.0)   # activ_density, base.yaml:98
    prog.gemm(w_last[1:], accumulate=False)
    n_feat = last.out_features - 1

    def feature_epilogue():
        prog.epilogue(last.bias[1:], hidden, n_feat, to_a=True)
        return n_feat

    bias, n_out = _append_mlp(prog, rgb_mlp, view_enc, _act_name(rgb_mlp.activ), first_epilogue=feature_epilogue)
    rgb = torch.empty(m, n_out, dtype=torch.float32, device=points_enc.device)
    prog.epilogue(bias, "sigmoid", n_out, to_a=False, out=rgb)
    prog.end()
    prog.run(m)
    return rgb, density[:, 0]


class _Truncated:
    """View of an MLPwithSkipConnection without its last layer."""

    def __init__(self, mlp):
        self.linears = list(mlp.linears)[:-1]
        self.skip_connection = mlp.skip_connection
        self.activ = mlp.activ


class FusedMlpFunction(torch.autograd.Function):
    """Differentiable MLPwithSkipConnection: fused tensor-core forward that also stores each hidden activation;
    backward with library GEMMs on the stored activations."""

    @staticmethod
    def forward(ctx, x, mlp, final_act, *params):
        layers = list(mlp.linears)
        saves = [torch.empty(x.shape[0], l.out_features, dtype=torch.float32, device=x.device) for l in layers[:-1]]
        with torch.no_grad():
            out = mlp_forward(mlp, x, final_act=final_act, saves=saves)
        ctx.mlp, ctx.final_act = mlp, final_act
        ctx.save_for_backward(x, out, *saves)
        return out

    @staticmethod
    def backward(ctx, d_out):
        x, out, *saves = ctx.saved_tensors
        mlp = ctx.mlp
        layers = list(mlp.linears)
        with torch.enable_grad():
            ws = [l.effective_weight() for l in layers]           # graph to (weight_g, weight_v) for the final chain rule
        d = d_out.float()
        if ctx.final_act == "sigmoid":
            d = d * out * (1.0 - out)
        grads_w, grads_b = [None] * len(layers), [None] * len(layers)
        dx = torch.zeros_like(x) if ctx.needs_input_grad[0] else None
        k0 = x.shape[1]
        for li in range(len(layers) - 1, -1, -1):
            a_in = x if li == 0 else saves[li - 1]
            if li in mlp.skip_connection and li > 0:
                a_in = torch.cat([a_in, x], dim=-1)
            grads_w[li] = d.t() @ a_in
            grads_b[li] = d.sum(dim=0)
            if li == 0:
                if dx is not None:
                    dx += d @ ws[0].detach()
                break
            da = d @ ws[li].detach()
            if li in mlp.skip_connection:
                if dx is not None:
                    dx += da[:, -k0:]
                da = da[:, :-k0]
            d = da * (saves[li - 1] > 0)                      # relu'
        # chain rule through the weight-norm reparametrisation
        flat_params = [p for l in layers for p in l.parameters()]
        wn_grads = torch.autograd.grad(ws, [p for l in layers for p in _weight_params(l)], grads_w, allow_unused=True)
        out_grads, it = [], iter(wn_grads)
        for l, gb in zip(layers, grads_b):
            for name, p in l.named_parameters():
                out_grads.append(gb if name == "bias" else next(it))
        assert len(out_grads) == len(flat_params)
        return (dx, None, None, *out_grads)


def _weight_params(lin):
    return [p for n, p in lin.named_parameters() if n != "bias"]


def fused_mlp(mlp, x, final_act="none"):
    """Autograd-aware fused evaluation of an MLPwithSkipConnection."""
    params = [p for l in mlp.linears for p in l.parameters()]
    return FusedMlpFunction.apply(x, mlp, final_act, *params)
