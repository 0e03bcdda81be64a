"""Host side of the tcgen05 fused SDF MLP (csrc/mlp_tcgen05.cu): weight packing and the forward call.

`pack_weight` turns an effective weight matrix W [N, K] (weight norm already applied) into the two fp16
images the kernel streams into shared memory: W = W_hi + 2^-11 W_lo', K zero-padded, laid out
[K/8][N/8][8][8] (K-chunk-major core matrices).  Packing costs one pass over ~0.1 M weights and is redone
only when the weights change (once per optimiser step), not per call.
"""
import torch

from . import _lib
from ._lib import ptr

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
        with torch.no_grad():
            self.w0_hi, self.w0_lo = pack_weight(l0.effective_weight(), self.k0_pad)
            self.w1_hi, self.w1_lo = pack_weight(l1.effective_weight(), self.n)
            self.b0 = l0.bias.detach().float().contiguous()
            self.b1 = l1.bias.detach().float().contiguous()
            self.w_head = mlp.linear_sdf.weight.detach().float().reshape(-1).contiguous()
            self.b_head = float(mlp.linear_sdf.bias.detach().float().item())

    def forward(self, rows, with_sdf=True, with_feat=True, with_hidden=False):
        """rows: [M, K0] fp32 CUDA tensor, unit stride in the last dim (row stride free).
        Returns (sdf [M,1] | None, feat [M,N] | None[, hidden [M,N]])."""
        _lib.require_cuda(rows)
        assert rows.dim() == 2 and rows.shape[1] == self.k0 and rows.stride(1) == 1 and rows.dtype == torch.float32
        m = rows.shape[0]
        dev = rows.device
        sdf = torch.empty(m, 1, dtype=torch.float32, device=dev) if with_sdf else None
        feat = torch.empty(m, self.n, dtype=torch.float32, device=dev) if with_feat else None
        hidden = torch.empty(m, self.n, dtype=torch.float32, device=dev) if with_hidden else None
        rc = _L().na_sdf_mlp_fwd(ptr(rows), rows.stride(0), m, self.k0, self.n, 2 if with_feat else 1,
                                 ptr(self.w0_hi), ptr(self.w0_lo), ptr(self.b0), self.k0_pad,
                                 ptr(self.w1_hi), ptr(self.w1_lo), ptr(self.b1),
                                 ptr(self.w_head) if with_sdf else None, self.b_head, self.beta,
                                 ptr(sdf), ptr(feat), self.n, ptr(hidden), self.n, _lib.current_stream())
        _lib.check(rc, "na_sdf_mlp_fwd")
        return (sdf, feat, hidden) if with_hidden else (sdf, feat)
