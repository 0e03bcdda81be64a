"""Host side of the generic tcgen05 MLP chain (csrc/mlp_chain.cu): builds the stage programs for
`MLPwithSkipConnection` (colour network) and for the whole `BackgroundNeRF` (feature MLP -> density + view-dependent
colour MLP in ONE launch), and wraps them for autograd.

Forward runs fused on the tensor cores with the split-fp16 arithmetic; for training the forward also saves each
layer's activations, the backward's input-gradient chain is a second stage program on the same kernel (relu-mask
epilogues reading the saved activations) and the weight / bias gradients come from the tcgen05 weight-gradient kernel
(csrc/wgrad_tcgen05.cu).
"""
import ctypes

import torch

from . import _lib
from ._lib import MlpStage, ptr
from .fused_mlp import pack_weight, pad16

_L = _lib.lib
ACT = {"none": 0, "relu": 1, "softplus": 2, "sigmoid": 3, "relu_mask": 4}
MAX_K = 256


# ---- packed-operand cache ---------------------------------------------------------------------------------------------
_PACKS = {}          # (id(lin), rows, cols, transposed) -> entry
_EPOCH = [0]


def _block(lin, rows, cols, transposed):
    with torch.no_grad():
        w = lin.effective_weight().detach().float()[rows[0]:rows[1], cols[0]:cols[1]]
        w = w.t() if transposed else w
        n_out, k = w.shape
        wp = torch.zeros(pad16(n_out), pad16(k), dtype=torch.float32, device=w.device)
        wp[:n_out, :k] = w
        return wp


def _packed_block(lin, rows, cols, transposed):
    key = (id(lin), tuple(rows), tuple(cols), bool(transposed))
    versions = tuple(p._version for p in lin.parameters())
    ent = _PACKS.get(key)
    if ent is None or ent["lin"]() is not lin or ent["epoch"] != _EPOCH[0] or ent["versions"] != versions:
        wp = _block(lin, rows, cols, transposed)
        hi, lo = pack_weight(wp, wp.shape[1])
        if ent is not None and ent["lin"]() is lin and ent["hi"].shape == hi.shape and ent["hi"].device == hi.device:
            ent["hi"].copy_(hi), ent["lo"].copy_(lo)         # in place: addresses baked into a captured graph stay valid
        else:
            import weakref
            ent = dict(lin=weakref.ref(lin), hi=hi, lo=lo, n_pad=wp.shape[0], k_pad=wp.shape[1], rows=tuple(rows),
                       cols=tuple(cols), transposed=bool(transposed))
            _PACKS[key] = ent
        ent["epoch"], ent["versions"] = _EPOCH[0], versions
    return ent["hi"], ent["lo"], ent["n_pad"], ent["k_pad"]


def invalidate_packs():
    """The parameters changed behind torch's back: cached operands are re-packed lazily at their next use."""
    _EPOCH[0] += 1


def refresh_packs():
    """Call after the parameters changed behind torch's back (the fused optimizer kernels): every cached operand is
    re-packed now, in place -- eagerly, i.e. outside any CUDA graph that uses it."""
    _EPOCH[0] += 1
    for key, ent in list(_PACKS.items()):
        lin = ent["lin"]()
        if lin is None:
            del _PACKS[key]
            continue
        _packed_block(lin, ent["rows"], ent["cols"], ent["transposed"])


class _Program:
    """Accumulates stages; keeps every tensor whose address is baked into the program alive."""

    def __init__(self):
        self.stages, self.keep = [], []
        self._cur = MlpStage()

    def epilogue(self, bias, act, n_valid, to_a, out=None, save=None, act_param=0.0):
        """act "relu_mask": `save` is READ (the layer's saved activation) and the accumulator is passed where it is > 0."""
        s = self._cur
        if bias is None:
            bias = torch.zeros(n_valid, dtype=torch.float32, device=(out if out is not None else save).device)
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

    def gemm(self, lin, rows, cols, accumulate, transposed=False):
        """Product with the block W[rows[0]:rows[1], cols[0]:cols[1]] of `lin`'s effective weight (its transpose when
        `transposed`): the block is [n_out, k], k multiplying the columns currently staged in A.  The packed operand
        comes from a per-layer cache that is refreshed in place once per optimizer step (refresh_packs), so that a
        training step -- and a CUDA graph of it -- packs nothing."""
        hi, lo, n_pad, k_pad = _packed_block(lin, rows, cols, transposed)
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
        allr = (0, lin.out_features)
        if li == 0:
            if first_epilogue is None:
                for c, n in _input_segments(k0):
                    prog.load(x, c, n)
                    prog.gemm(lin, allr, (c, c + n), accumulate=c > 0)
            else:
                resident = first_epilogue()            # emits the epilogue that fills A; returns its width
                prog.gemm(lin, allr, (0, resident), accumulate=False)
                for c, n in _input_segments(k0):
                    prog.load(x, c, n)
                    prog.gemm(lin, allr, (resident + c, resident + c + n), accumulate=True)
            continue
        prev = layers[li - 1]
        n_prev = prev.out_features
        assert n_prev <= MAX_K, "hidden width > 256 is outside the fused kernel's envelope"
        prog.epilogue(prev.bias, hidden_act, n_prev, to_a=True, save=None if saves is None else saves[li - 1])
        prog.gemm(lin, allr, (0, n_prev), accumulate=False)
        if li in mlp.skip_connection:      # feat = cat([feat, input]) before this layer (nerf_util.py:189-190)
            assert first_epilogue is None
            for c, n in _input_segments(k0):
                prog.load(x, c, n)
                prog.gemm(lin, allr, (n_prev + c, n_prev + c + n), accumulate=True)
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
    layers = list(mlp.linears)
    return all(l.out_features <= MAX_K for l in layers[:-1]) and layers[-1].out_features <= 2 * MAX_K


def mlp_forward(mlp, x, final_act="none", saves=None):
    """Fused forward of an MLPwithSkipConnection on rows x [M,K0] (fp32, unit inner stride) -> [M, n_out]."""
    _lib.require_cuda(x)
    m = x.shape[0]
    prog = _Program()
    layers = list(mlp.linears)
    n_out = layers[-1].out_features
    if n_out <= MAX_K:
        bias, n_out = _append_mlp(prog, mlp, x, _act_name(mlp.activ), saves=saves)
        out = torch.empty(m, n_out, dtype=torch.float32, device=x.device)
        prog.epilogue(bias, final_act, n_out, to_a=False, out=out)
    else:
        # a last layer wider than one accumulator (the background feature MLP's 257 columns): every <= 256-column group
        # of its outputs is a product of the same A operand (the previous layer's activation)
        assert len(layers) >= 2 and (len(layers) - 1) not in mlp.skip_connection
        _append_mlp(prog, _Truncated(mlp), x, _act_name(mlp.activ), saves=saves)
        prev, last = layers[-2], layers[-1]
        out = torch.empty(m, n_out, dtype=torch.float32, device=x.device)
        prog.epilogue(prev.bias, _act_name(mlp.activ), prev.out_features, to_a=True,
                      save=None if saves is None else saves[len(layers) - 2])
        for c in range(0, n_out, MAX_K):
            n = min(MAX_K, n_out - c)
            prog.gemm(last, (c, c + n), (0, last.in_features), accumulate=False)
            prog.epilogue(last.bias[c:c + n], final_act, n, to_a=False, out=out[:, c:c + n])
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
    prev = feat_mlp.linears[-2]
    # last layer of mlp_feat as two products of the same A operand: column 0 (density), then columns 1.. (feature)
    prog.epilogue(prev.bias, hidden, prev.out_features, to_a=True)
    prog.gemm(last, (0, 1), (0, last.in_features), accumulate=False)
    density = torch.empty(m, 1, dtype=torch.float32, device=points_enc.device)
    prog.epilogue(last.bias[:1], "softplus", 1, to_a=False, out=density, act_param=1.0)   # activ_density, base.yaml:98
    prog.gemm(last, (1, last.out_features), (0, last.in_features), accumulate=False)
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


def mlp_backward(mlp, d_last, saves, k0, need_dx):
    """Input-gradient chain of an MLPwithSkipConnection as ONE stage program on the tensor cores:
        dz_L = d_last (the caller has applied the last activation's derivative and a power-of-two scale),
        dz_{l-1} = (dz_l W_l)[:, :n_{l-1}] * step(a_{l-1})          (relu'; a_{l-1} = saved activation),
        dx = dz_0 W_0 (+ the skip layers' input columns)            (only when need_dx).
    Returns (dz list, one [M, n_l] tensor per layer -- what the weight-gradient kernel reduces -- and dx or None)."""
    layers = list(mlp.linears)
    assert len(layers) >= 2, "the chain programs cover MLPs of at least two layers"
    m, dev = d_last.shape[0], d_last.device
    skips = [li for li in mlp.skip_connection if 0 < li < len(layers)]
    if need_dx and skips:
        raise NotImplementedError("input gradient through a skip connection (not needed on the path: the skip networks "
                                  "take encodings of non-differentiable points)")
    dz = [None] * len(layers)
    dz[-1] = d_last
    prog = _Program()
    n_last = layers[-1].out_features
    for li in range(len(layers) - 1, 0, -1):
        n_prev = layers[li - 1].out_features
        lin = layers[li]                                  # product D = dz_l W_l[:, :n_prev]: operand (W_l[:, :n_prev])^T
        if li == len(layers) - 1:
            for c, n in _input_segments(n_last):          # dz_L arrives from global memory, <= 256 columns at a time
                prog.load(d_last, c, n)
                prog.gemm(lin, (c, c + n), (0, n_prev), accumulate=c > 0, transposed=True)
        else:                                             # dz_l is resident in A (left there by the previous epilogue)
            prog.gemm(lin, (0, lin.out_features), (0, n_prev), accumulate=False, transposed=True)
        dz[li - 1] = torch.empty(m, n_prev, dtype=torch.float32, device=dev)
        prog.epilogue(None, "relu_mask", n_prev, to_a=True, out=dz[li - 1], save=saves[li - 1])
    dx = None
    if need_dx:
        dx = torch.empty(m, k0, dtype=torch.float32, device=dev)
        for c in range(0, k0, MAX_K):                     # dx in <= 256-column groups, all products of the same A = dz_0
            n = min(MAX_K, k0 - c)
            prog.gemm(layers[0], (0, layers[0].out_features), (c, c + n), accumulate=False, transposed=True)
            prog.epilogue(None, "none", n, to_a=False, out=dx[:, c:c + n])
    prog.end()
    prog.run(m)
    return dz, dx


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
        """Input-gradient chain on the tcgen05 chain kernel (mlp_backward), weight / bias gradients on the tcgen05
        weight-gradient kernel (mlp.linear_wgrad); torch only applies the final chain rule through weight norm."""
        from .mlp import linear_wgrad
        x, out, *saves = ctx.saved_tensors
        mlp = ctx.mlp
        layers = list(mlp.linears)
        with torch.enable_grad():
            ws = [l.effective_weight() for l in layers]           # graph to (weight_g, weight_v) for the final chain rule
        d = d_out.float()
        if ctx.final_act == "sigmoid":
            d = d * out * (1.0 - out)
        # gradients are tiny (1e-9 ... 1e-4): lift them by a power of two so that the fp16 pairs of the chain kernel keep
        # their mantissa; everything downstream is linear in d, so the factor is divided out at the end
        amax = torch.linalg.vector_norm(d, ord=float("inf")).clamp_min(1e-30)
        scale = torch.exp2(torch.floor(torch.log2(1024.0 / amax)))
        inv = 1.0 / scale
        d = (d * scale).contiguous()
        k0 = x.shape[1]
        dz, dx = mlp_backward(mlp, d, saves, k0, need_dx=ctx.needs_input_grad[0])
        if dx is not None:
            dx = dx * inv
        grads_w, grads_b = [None] * len(layers), [None] * len(layers)
        for li in range(len(layers)):
            a_in = x if li == 0 else saves[li - 1]
            if li in mlp.skip_connection and li > 0:
                a_in = torch.cat([a_in, x], dim=-1)
            gw, gb = linear_wgrad(dz[li], a_in)
            grads_w[li], grads_b[li] = gw * inv, gb * inv
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
