"""CPU: host-side math helpers of the training path (pure torch, so they run here): the split-K weight-gradient product,
the blocked column sum and the Linear autograd function built on them must agree with the plain expressions."""
import torch
import torch.nn.functional as F


def test_tn_matmul_and_colsum_match_the_plain_expressions():
    from neuralangelo_b200.mlp import colsum, tn_matmul
    g = torch.Generator().manual_seed(0)
    for m in (8192, 65536, 1000, 4096 + 64):      # split into 8 / 64 parts, and two shapes that fall back
        a = torch.randn(m, 24, generator=g, dtype=torch.float64)
        b = torch.randn(m, 19, generator=g, dtype=torch.float64)
        torch.testing.assert_close(tn_matmul(a, b), a.t() @ b, rtol=1e-12, atol=1e-10)
        torch.testing.assert_close(colsum(a), a.sum(dim=0), rtol=1e-12, atol=1e-10)
    # strided right operand (the [M,131] row view of a 132-wide buffer)
    buf = torch.randn(8192, 20, generator=g, dtype=torch.float64)
    a = torch.randn(8192, 7, generator=g, dtype=torch.float64)
    torch.testing.assert_close(tn_matmul(a, buf[:, 1:]), a.t() @ buf[:, 1:], rtol=1e-12, atol=1e-10)
    # non-contiguous input of colsum takes the fallback
    torch.testing.assert_close(colsum(buf[:, 1:]), buf[:, 1:].sum(dim=0), rtol=1e-12, atol=1e-10)


def test_linear_tall_rows_function_has_f_linear_gradients():
    from neuralangelo_b200.mlp import _LinearTallRows
    g = torch.Generator().manual_seed(1)
    x = torch.randn(8192, 13, generator=g, dtype=torch.float64, requires_grad=True)
    w = torch.randn(5, 13, generator=g, dtype=torch.float64, requires_grad=True)
    b = torch.randn(5, generator=g, dtype=torch.float64, requires_grad=True)
    dy = torch.randn(8192, 5, generator=g, dtype=torch.float64)
    ref = torch.autograd.grad(F.linear(x, w, b), [x, w, b], dy)
    got = torch.autograd.grad(_LinearTallRows.apply(x, w, b), [x, w, b], dy)
    for r, k in zip(ref, got):
        torch.testing.assert_close(k, r, rtol=1e-12, atol=1e-10)
    torch.testing.assert_close(_LinearTallRows.apply(x, w, b), F.linear(x, w, b), rtol=1e-12, atol=1e-12)


def test_linear_dispatch_is_f_linear_on_the_cpu():
    """The CPU path (oracle comparisons, golden fixtures) must stay the reference's own arithmetic."""
    from neuralangelo_b200 import mlp
    lin = mlp.Linear(7, 4, weight_norm=True)
    x = torch.randn(9000, 7)
    w = lin.weight_v * (lin.weight_g / lin.weight_v.norm(dim=1, keepdim=True))
    assert torch.equal(lin(x), F.linear(x, w, lin.bias))
