"""Which library GEMM does the colour network's forward get?  addmm [M,K]x[K,256]+b for M = 65536 in one call / in row chunks."""
import torch
torch.backends.cuda.matmul.allow_tf32 = True
dev = torch.device("cuda:0")
def t(fn, n=20):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for K, ld in ((256, 256), (278, 278), (278, 280), (280, 280)):
    M = 65536
    xf = torch.randn(M, ld, device=dev)
    x = xf[:, :K]
    wf = torch.randn(256, ld, device=dev)
    w = wf[:, :K]
    b = torch.randn(256, device=dev)
    y = torch.empty(M, 256, device=dev)
    res = {"K": K, "ld": ld}
    res["addmm_full"] = t(lambda: torch.addmm(b, x, w.t(), out=y))
    for c in (32768, 16384, 8192):
        res[f"addmm_chunks_{c}"] = t(lambda: [torch.addmm(b, x[i:i + c], w.t(), out=y[i:i + c]) for i in range(0, M, c)])
    res["mm_full"] = t(lambda: torch.mm(x, w.t(), out=y))
    dy = torch.randn(M, 256, device=dev)
    dxf = torch.empty(M, ld, device=dev)
    res["dx_mm_full"] = t(lambda: torch.mm(dy, w, out=dxf[:, :K]) if ld != K else torch.mm(dy, w))
    res["dx_chunks_16384"] = t(lambda: [torch.mm(dy[i:i + 16384], w, out=dxf[i:i + 16384, :K]) for i in range(0, M, 16384)])
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in res.items()})
