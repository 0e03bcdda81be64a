"""CPU, world_size 2 over gloo: the host-side multi-GPU logic (gradient bucket all-reduce with the
active-level prefix, block sharding, variable-length mesh gather)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, ret), nprocs=world, join=True)
    return [ret[r] for r in range(world)]


def _allreduce_case(rank, world):
    from neuralangelo_b200 import parallel
    flat = torch.full((10,), float(rank + 1))
    table = torch.arange(20, dtype=torch.float32) * (rank + 1)
    prefix = table[:12]                                   # "active levels" prefix; the tail must stay untouched
    scale = parallel.allreduce_gradients(flat, prefix)
    return flat.tolist(), table.tolist(), scale


def test_gradient_allreduce_sums_bucket_and_active_prefix_only():
    out = _run(_allreduce_case)
    for rank, (flat, table, scale) in enumerate(out):
        assert scale == 0.5
        assert flat == [3.0] * 10                          # 1 + 2
        expect = [i * 3.0 for i in range(12)] + [i * float(rank + 1) for i in range(12, 20)]
        assert table == expect


def _gather_case(rank, world):
    from neuralangelo_b200 import mesh
    n = 3 if rank == 0 else 5
    v = torch.full((n, 3), float(rank), dtype=torch.float64)
    f = torch.arange(n * 3, dtype=torch.long).reshape(n, 3) + 100 * rank
    vs, sizes = mesh._gather_to_rank0(v, rank, world, torch.device("cpu"))
    fs, _ = mesh._gather_to_rank0(f, rank, world, torch.device("cpu"))
    empty, esizes = mesh._gather_to_rank0(torch.zeros(0, 3, dtype=torch.float64), rank, world, torch.device("cpu"))
    if rank != 0:
        return vs is None and fs is None and empty is None, sizes
    return [t.tolist() for t in vs], [t.tolist() for t in fs], sizes, [len(e) for e in empty]


def test_mesh_gather_variable_lengths():
    out = _run(_gather_case)
    vs, fs, sizes, empty = out[0]
    assert sizes == [3, 5] and empty == [0, 0]
    assert vs[0] == [[0.0] * 3] * 3 and vs[1] == [[1.0] * 3] * 5
    assert fs[1][0] == [100, 101, 102]
    assert out[1] == (True, [3, 5])


def test_block_sharding_is_round_robin_without_padding():
    from neuralangelo_b200 import parallel
    shards = [parallel.shard_indices(10, r, 4) for r in range(4)]
    assert shards == [[0, 4, 8], [1, 5, 9], [2, 6], [3, 7]]
    assert sorted(sum(shards, [])) == list(range(10))
    # 4096 blocks (res 2048 / block 128) over 8 ranks: 512 each, block idx = rank (mod 8), as DistributedSampler(shuffle=False)
    s8 = [parallel.shard_indices(4096, r, 8) for r in range(8)]
    assert all(len(s) == 512 for s in s8) and s8[3][:3] == [3, 11, 19]


def test_flat_bucket_puts_every_parameter_on_a_256_byte_boundary():
    """FusedAdamW's bucket: each small parameter (and its gradient view) starts on a multiple of 64 floats -- the library
    GEMMs that still read biases / weights on the training path pick their vectorised kernels only for 16-byte aligned
    operands -- the views alias the bucket, and the padding between parameters is zero."""
    import torch
    from neuralangelo_b200.optim import FusedAdamW
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Linear(5, 3), torch.nn.Linear(3, 257))
    before = [p.detach().clone() for p in model.parameters()]
    opt = FusedAdamW(model)
    base_p, base_g = opt.flat_p.data_ptr(), opt.flat_g.data_ptr()
    covered = torch.zeros(opt.flat_p.numel(), dtype=torch.bool)
    for p, ref, g in zip(model.parameters(), before, opt._grad_views):
        off = (p.data_ptr() - base_p) // 4
        assert off % 64 == 0 and (g.data_ptr() - base_g) // 4 == off
        assert torch.equal(p.detach(), ref)                      # values moved into the bucket unchanged
        covered[off:off + p.numel()] = True
    assert opt.flat_p.numel() % 64 == 0 and float(opt.flat_p[~covered].abs().max()) == 0.0
    model[0].bias.data.fill_(3.0)                                # the parameter IS the bucket
    off = (model[0].bias.data_ptr() - base_p) // 4
    assert torch.equal(opt.flat_p[off:off + 5], torch.full((5,), 3.0))
