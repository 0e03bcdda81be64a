"""Exchange-only microbenchmark (torchrun, one rank per GPU): the fused multicast kernel (reduce + AdamW + broadcast of
the full default table, 365.8 M fp32) for several grid sizes, against NCCL reduce-scatter + AdamW slice + all-gather.
CUDA events on each rank, max over ranks, median of 10."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
from neuralangelo_b200 import _lib, build, parallel

rank, local, world = parallel.init_dist()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
build.build(verbose=False)
L = _lib.lib()
n = 45724048 * 8
p = symm_mem.empty(n, dtype=torch.float32, device=dev)
g = symm_mem.empty(n, dtype=torch.float32, device=dev)
hp, hg = symm_mem.rendezvous(p, dist.group.WORLD), symm_mem.rendezvous(g, dist.group.WORLD)
p.normal_(); g.normal_()
m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
part = n // world // 4 * 4
stream = _lib.current_stream()


def timed(fn, iters=10, warm=3):
    ts = []
    for i in range(warm + iters):
        dist.barrier(); torch.cuda.synchronize()
        s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); t.record(); torch.cuda.synchronize()
        if i >= warm:
            ts.append(s.elapsed_time(t))
    x = torch.tensor([float(np.median(ts))], device=dev, dtype=torch.float64)
    dist.all_reduce(x, op=dist.ReduceOp.MAX)
    return round(float(x), 4)


res = {"world": world, "multicast": bool(hp.multicast_ptr), "table_GB": n * 4 / 1e9}
for ctas, unroll in ((37, 4), (74, 4), (148, 2), (148, 4), (148, 8), (296, 8), (1184, 4)):
    def fused():
        hg.barrier(channel=0)
        _lib.check(L.na_adamw_step_multimem(ctypes.c_void_p(hg.multicast_ptr), ctypes.c_void_p(hp.multicast_ptr), _lib.ptr(p),
                                            _lib.ptr(m), _lib.ptr(v), rank * part, part, 1e-3, 0.9, 0.999, 1e-8, 1e-2, 1,
                                            1.0 / world, ctas, unroll, stream), "na_adamw_step_multimem")
        hp.barrier(channel=1)
        g[:part * world].zero_()
    res[f"fused_ms_ctas{ctas}_unroll{unroll}"] = timed(fused)


def kernel_only():
    _lib.check(L.na_adamw_step_multimem(ctypes.c_void_p(hg.multicast_ptr), ctypes.c_void_p(hp.multicast_ptr), _lib.ptr(p),
                                        _lib.ptr(m), _lib.ptr(v), rank * part, part, 1e-3, 0.9, 0.999, 1e-8, 1e-2, 1,
                                        1.0 / world, 148, 4, stream), "na_adamw_step_multimem")


res["fused_kernel_only_ms"] = timed(kernel_only)
res["two_barriers_ms"] = timed(lambda: (hg.barrier(channel=0), hp.barrier(channel=1)))
res["local_grad_zero_ms"] = timed(lambda: g[:part * world].zero_())
a, b = rank * part, (rank + 1) * part


def nccl():
    dist.reduce_scatter_tensor(g[a:b], g[:part * world], op=dist.ReduceOp.SUM)
    _lib.check(L.na_adamw_step(_lib.ptr(p[a:]), _lib.ptr(g[a:]), _lib.ptr(m[a:]), _lib.ptr(v[a:]), part, 1e-3, 0.9, 0.999,
                               1e-8, 1e-2, 1, 1.0 / world, 1, stream), "na_adamw_step")
    dist.all_gather_into_tensor(p[:part * world], p[a:b])


res["nccl_rs_adamw_ag_ms"] = timed(nccl)
res["single_gpu_adamw_full_table_ms"] = timed(lambda: _lib.check(L.na_adamw_step(_lib.ptr(p), _lib.ptr(g), _lib.ptr(m), _lib.ptr(v), n,
                                                                                 1e-3, 0.9, 0.999, 1e-8, 1e-2, 1, 1.0, 1, stream), "adamw"))
if rank == 0:
    print(json.dumps(res), flush=True)
dist.barrier()
dist.destroy_process_group()
