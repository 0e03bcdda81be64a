"""Host side of the tcgen05 fused SDF MLP (csrc/mlp_tcgen05.cu): weight packing and the forward call.

`pack_weight` turns an effective weight matrix W [N, K] (weight norm already applied) into the two fp16
images the kernel streams into shared memory: W = W_hi + 2^-11 W_lo', K zero-padded, laid out
[K/8][N/8][8][8] (K-chunk-major core matrices).  Packing costs one pass over ~0.1 M weights and is redone
only when the weights change (once per optimiser step), not per call.
"""
import math

import torch

from . import _lib
from ._lib import ptr
from .mlp import colsum, linear_wgrad, tn_matmul, wgrad_supported

_L = _lib.lib
LO_SCALE = 2048.0


def pad16(k):
    return (k + 15) // 16 * 16


def pack_weight(w, k_pad):
    """W [N,K] fp32 -> (hi, lo) fp16 tensors of N*k_pad elements in the kernel's layout."""
    n, k = w.shape
    assert n % 8 == 0 and k_pad % 16 == 0 and k <= k_pad
    wp = torch.zeros(n, k_pad, dtype=torch.float32, device=w.device)
    wp[:, :k] = w.detach().float().clamp(-65504.0, 65504.0)
    hi = wp.half()
    lo = ((wp - hi.float()) * LO_SCALE).half()

    def lay(t):
        return t.view(n // 8, 8, k_pad // 8, 8).permute(2, 0, 1, 3).contiguous().view(-1)

    return lay(hi), lay(lo)


class PackedSdfMlp:
    """Packed weights of an MLPforNeuralSDF with one hidden layer (base.yaml: num_layers 1, no skip)."""

    def __init__(self, mlp, beta):
        if len(mlp.linears) != 2 or mlp.skip_connection:
            raise NotImplementedError("fused SDF MLP covers the shipped architecture: num_layers=1, no skip connection")
        l0, l1 = mlp.linears
        self.n = l0.out_features
        if self.n % 64 != 0 or self.n > 256 or l1.out_features != self.n:
            raise NotImplementedError(f"hidden width {self.n} unsupported by the fused kernel (64/128/192/256)")
        self.k0 = l0.in_features
        self.k0_pad = pad16(self.k0)
        if self.k0_pad > 256:
            raise NotImplementedError("input width > 256 unsupported by the fused kernel")
        self.beta = float(beta)
        self.update_(mlp)

    def _pack(self, name, w, n, k, k_pad, n_valid=None, transposed=False, rot=0, scale=1.0, extra=None):
        """One launch (na_pack_split_f16) into the persistent <name>_hi / <name>_lo tensors: kernels recorded in a CUDA
        graph keep reading valid, current data from the same addresses."""
        hi = getattr(self, name + "_hi", None)
        if hi is None or hi.numel() != n * k_pad or hi.device != w.device:
            hi = torch.empty(n * k_pad, dtype=torch.float16, device=w.device)
            setattr(self, name + "_hi", hi)
            setattr(self, name + "_lo", torch.empty_like(hi))
        assert w.dtype == torch.float32 and w.stride(1) == 1
        rc = _L().na_pack_split_f16(ptr(w), w.stride(0), n, n if n_valid is None else n_valid, k, int(transposed), rot,
                                    float(scale), ptr(extra), k_pad, ptr(hi), ptr(getattr(self, name + "_lo")),
                                    _lib.current_stream())
        _lib.check(rc, "na_pack_split_f16")

    def update_(self, mlp):
        """(Re)pack from the module's current weights, in place after the first call."""
        l0, l1 = mlp.linears
        n, k0, k0_pad = self.n, self.k0, self.k0_pad
        with torch.no_grad():
            w0 = l0.effective_weight().detach().float().contiguous()
            w1 = l1.effective_weight().detach().float().contiguous()
            new = {"b0": l0.bias.detach().float().contiguous(), "b1": l1.bias.detach().float().contiguous(),
                   "w_head": mlp.linear_sdf.weight.detach().float().reshape(-1).contiguous(),
                   "b_head": mlp.linear_sdf.bias.detach().float().reshape(1),
                   # the two norms the backward's row scaling uses
                   "norms": torch.stack([w1.abs().sum(dim=0).max(), mlp.linear_sdf.weight.detach().abs().max().float()])}
            for name, t in new.items():
                cur = getattr(self, name, None)
                if cur is not None and cur.shape == t.shape and cur.device == t.device:
                    cur.copy_(t)
                else:
                    setattr(self, name, t.clone())
            self._pack("w0", w0, n, k0, k0_pad)
            self._pack("w1", w1, n, n, n)
            # chunk-plane column order: encoding features first (one 8-wide chunk per level), xyz in the last chunk
            self._pack("w0p", w0, n, k0, k0_pad, rot=3)
            # operands of the backward products (dz1 W1, dz0 W0): the transposed weights
            self._pack("w0t", w0, k0_pad, n, n, n_valid=k0, transposed=True)
            self._pack("w1t", w1, n, n, n, transposed=True)
            if self.ws_supported():
                # operands of the warp-specialised SDF-only kernels: c * [W0[:, 3:] | W0[:, :3] | b0 | 0 ...] with
                # c = beta * log2(e) -- the bias rides on a constant-one input column, the accumulator is the softplus
                # argument in the log2 domain -- and w_head * ln(2) / beta
                self._pack("w0s", w0, n, k0, k0_pad, rot=3, scale=self.beta * math.log2(math.e), extra=self.b0)
                head = self.w_head * (math.log(2.0) / self.beta)
                if getattr(self, "w_head_s", None) is not None and self.w_head_s.device == head.device:
                    self.w_head_s.copy_(head)
                else:
                    self.w_head_s = head

    def lattice_operands(self):
        """(w0s_hi, w0s_lo, w_head_s): the pre-scaled operands of na_lattice_sdf / na_sdf_head_ws (see update_)."""
        if not self.ws_supported():
            raise NotImplementedError("warp-specialised SDF-only kernels cover base.yaml's network (131 -> 256 -> 1)")
        return self.w0s_hi, self.w0s_lo, self.w_head_s

    def ws_supported(self):
        """Shape of the warp-specialised SDF-only kernels (na_lattice_sdf / na_sdf_head_ws): base.yaml's network."""
        return self.n == 256 and self.k0 == 131 and self.k0_pad == 144

    def _head_ws(self, planes, chunk_stride, rows, m):
        w_hi, w_lo, head = self.lattice_operands()
        dev = planes.device if planes is not None else rows.device
        sdf = torch.empty(m, 1, dtype=torch.float32, device=dev)
        rc = _L().na_sdf_head_ws(ptr(planes), chunk_stride, ptr(rows), 0 if rows is None else rows.stride(0), m, ptr(w_hi),
                                 ptr(w_lo), self.k0_pad, self.n, ptr(head), ptr(self.b_head), 1.0, ptr(sdf),
                                 _lib.current_stream())
        _lib.check(rc, "na_sdf_head_ws")
        return sdf

    def forward(self, rows, with_sdf=True, with_feat=True, with_hidden=False):
        """rows: [M, K0] fp32 CUDA tensor, unit stride in the last dim (row stride free).
        Returns (sdf [M,1] | None, feat [M,N] | None[, hidden [M,N]])."""
        _lib.require_cuda(rows)
        assert rows.dim() == 2 and rows.shape[1] == self.k0 and rows.stride(1) == 1 and rows.dtype == torch.float32
        m = rows.shape[0]
        dev = rows.device
        if with_sdf and not with_feat and not with_hidden and self.ws_supported():
            return self._head_ws(None, 0, rows, m), None
        sdf = torch.empty(m, 1, dtype=torch.float32, device=dev) if with_sdf else None
        feat = torch.empty(m, self.n, dtype=torch.float32, device=dev) if with_feat else None
        hidden = torch.empty(m, self.n, dtype=torch.float32, device=dev) if with_hidden else None
        rc = _L().na_sdf_mlp_fwd(ptr(rows), rows.stride(0), m, self.k0, self.n, 2 if with_feat else 1,
                                 ptr(self.w0_hi), ptr(self.w0_lo), ptr(self.b0), self.k0_pad,
                                 ptr(self.w1_hi), ptr(self.w1_lo), ptr(self.b1),
                                 ptr(self.w_head) if with_sdf else None, ptr(self.b_head), self.beta,
                                 ptr(sdf), ptr(feat), self.n, ptr(hidden), self.n, _lib.current_stream())
        _lib.check(rc, "na_sdf_mlp_fwd")
        return (sdf, feat, hidden) if with_hidden else (sdf, feat)


    def forward_planes(self, planes, m, with_hidden=False):
        """SDF-only evaluation of chunk planes: planes [C, M_alloc, 8] fp32 (C = ceil(k0 / 8) chunks; chunk c < C-1 is
        encoding level c, the last chunk holds xyz in its first three floats), rows [0, m).  Returns sdf [m,1]
        (and the hidden activations [m,n])."""
        _lib.require_cuda(planes)
        assert planes.dim() == 3 and planes.shape[2] == 8 and planes.stride(2) == 1 and planes.stride(1) == 8
        assert planes.shape[0] * 8 >= self.k0 and planes.shape[1] >= m and planes.dtype == torch.float32
        dev = planes.device
        if not with_hidden and self.ws_supported() and planes.shape[0] == 17:
            return self._head_ws(planes, planes.stride(0), None, m)
        sdf = torch.empty(m, 1, dtype=torch.float32, device=dev)
        hidden = torch.empty(m, self.n, dtype=torch.float32, device=dev) if with_hidden else None
        rc = _L().na_sdf_head_fwd_planes(ptr(planes), planes.stride(0), m, self.k0, self.n, ptr(self.w0p_hi),
                                         ptr(self.w0p_lo), ptr(self.b0), self.k0_pad, ptr(self.w_head), ptr(self.b_head),
                                         self.beta, ptr(sdf), ptr(hidden), self.n, _lib.current_stream())
        _lib.check(rc, "na_sdf_head_fwd_planes")
        return (sdf, hidden) if with_hidden else sdf


class FusedSdfMlpFunction(torch.autograd.Function):
    """Differentiable fused SDF MLP: forward and input-gradient chain on tcgen05; weight gradients
    (reductions over all rows) by library GEMMs on the pre-activation gradients the kernel returns.
    Arguments are the EFFECTIVE weights (torch applies / differentiates the weight-norm reparametrisation)."""

    @staticmethod
    def forward(ctx, rows, w0, b0, w1, b1, w_head, b_head, beta, with_feat, packed=None):
        """`packed`: a PackedSdfMlp holding the CURRENT weights (NeuralSDF keeps one and refreshes it after every
        optimiser step); without it the operands are packed here, per call."""
        m, k0 = rows.shape
        n = w0.shape[0]
        k0_pad = pad16(k0)
        if packed is not None:
            w0_hi, w0_lo, w1_hi, w1_lo = packed.w0_hi, packed.w0_lo, packed.w1_hi, packed.w1_lo
            b0c, b1c, whc, bhc = packed.b0, packed.b1, packed.w_head, packed.b_head
        else:
            with torch.no_grad():
                w0_hi, w0_lo = pack_weight(w0, k0_pad)
                w1_hi, w1_lo = pack_weight(w1, n) if with_feat else (None, None)
            b0c, b1c, whc = b0.detach().float().contiguous(), b1.detach().float().contiguous(), \
                w_head.detach().float().reshape(-1).contiguous()
            bhc = b_head.detach().float().reshape(1).contiguous()
        dev = rows.device
        sdf = torch.empty(m, 1, dtype=torch.float32, device=dev)
        feat = torch.empty(m, n, dtype=torch.float32, device=dev) if with_feat else None
        hidden = torch.empty(m, n, dtype=torch.float32, device=dev)
        rc = _L().na_sdf_mlp_fwd(ptr(rows), rows.stride(0), m, k0, n, 2 if with_feat else 1, ptr(w0_hi), ptr(w0_lo),
                                 ptr(b0c), k0_pad, ptr(w1_hi) if with_feat else None, ptr(w1_lo) if with_feat else None,
                                 ptr(b1c), ptr(whc), ptr(bhc), float(beta), ptr(sdf), ptr(feat), n, ptr(hidden), n,
                                 _lib.current_stream())
        _lib.check(rc, "na_sdf_mlp_fwd")
        ctx.save_for_backward(rows, hidden, feat if with_feat else hidden, w0, w1, whc)
        ctx.beta, ctx.with_feat, ctx.packed = float(beta), with_feat, packed
        if not with_feat:
            feat = torch.empty(0, device=dev)
            ctx.mark_non_differentiable(feat)
        return sdf, feat

    @staticmethod
    def backward(ctx, d_sdf, d_feat):
        rows, hidden, feat, w0, w1, w_head = ctx.saved_tensors
        m, k0 = rows.shape
        n = w0.shape[0]
        k0_pad = pad16(k0)
        dev = rows.device
        two = ctx.with_feat and d_feat is not None
        d_sdf = d_sdf.float().contiguous().reshape(-1) if d_sdf is not None else torch.zeros(m, device=dev)
        packed = ctx.packed
        with torch.no_grad():
            if packed is not None:
                w0t_hi, w0t_lo = packed.w0t_hi, packed.w0t_lo
            else:
                w0t = torch.zeros(k0_pad, n, dtype=torch.float32, device=dev)
                w0t[:k0] = w0.detach().float().t()
                w0t_hi, w0t_lo = pack_weight(w0t, n)
            if two:
                d_feat = d_feat.float().contiguous()
                w1t_hi, w1t_lo = (packed.w1t_hi, packed.w1t_lo) if packed is not None else \
                    pack_weight(w1.detach().float().t().contiguous(), n)
                bound = d_feat.abs().amax(dim=1).contiguous()
                w1_colsum = w1.detach().abs().sum(dim=0).max().reshape(1) if packed is None else None
                dz1 = torch.empty(m, n, dtype=torch.float32, device=dev)
            else:
                w1t_hi = w1t_lo = bound = dz1 = None
                w1_colsum = torch.zeros(1, device=dev) if packed is None else None
            if packed is not None:   # the two-layer bound is also valid (merely looser) for SDF-only rows
                norms = packed.norms
            else:
                norms = torch.cat([w1_colsum.float(), w_head.abs().max().reshape(1)])   # device scalars: no host sync
            dz0 = torch.empty(m, n, dtype=torch.float32, device=dev)
            dx = torch.empty(m, k0, dtype=torch.float32, device=dev)
            rc = _L().na_sdf_mlp_bwd(m, n, k0, k0_pad, 1 if two else 0, ptr(d_sdf), ptr(d_feat) if two else None,
                                     ptr(feat) if two else None, ptr(hidden), n, ptr(bound), ptr(w1t_hi), ptr(w1t_lo),
                                     ptr(w0t_hi), ptr(w0t_lo), ptr(w_head), ctx.beta, ptr(norms), ptr(dz1), ptr(dz0),
                                     ptr(dx), k0, _lib.current_stream())
            _lib.check(rc, "na_sdf_mlp_bwd")
            # weight gradients: reductions over the rows, in the ambient matmul precision (TrainStep enables TF32 by
            # default, as the reference's trainer does: imaginaire/trainers/base.py:172-175)
            if wgrad_supported(dz0, rows):      # tcgen05 weight-gradient kernel (csrc/wgrad_tcgen05.cu)
                dw0, db0 = linear_wgrad(dz0, rows)
                dw1, db1 = linear_wgrad(dz1, hidden) if two else (None, None)
                dwh, dbh = linear_wgrad(d_sdf[:, None], hidden)
            else:
                dw0 = tn_matmul(dz0, rows)
                db0 = colsum(dz0)
                dw1 = tn_matmul(dz1, hidden) if two else None
                db1 = colsum(dz1) if two else None
                dwh = (d_sdf[None, :] @ hidden)
                dbh = d_sdf.sum().reshape(1)
        return dx, dw0, db0, dw1, db1, dwh, dbh, None, None, None


def fused_sdf_mlp(mlp, rows, beta, with_sdf=True, with_feat=True, packed=None):
    """Autograd-aware fused evaluation of an MLPforNeuralSDF (one hidden layer) on rows [M,K0]."""
    l0, l1 = mlp.linears
    sdf, feat = FusedSdfMlpFunction.apply(rows, l0.effective_weight(), l0.bias, l1.effective_weight(), l1.bias,
                                          mlp.linear_sdf.weight, mlp.linear_sdf.bias, beta, with_feat, packed)
    return (sdf if with_sdf else None), (feat if with_feat else None)
