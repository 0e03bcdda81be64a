"""Data-parallel training-step timing only (the `extras.train_step*` entries of bench.py without the rest):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_dp.py
Rank 0 prints one JSON object: default config, 512 rays per GPU, 16 active levels, taps 4 and 6, fused multicast
exchange and the NCCL form."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from neuralangelo_b200 import build, parallel

rank, local, world = parallel.init_dist()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
build.build(verbose=False)
out = {"world": world}
for name, kw in (("taps4", dict(taps=4)), ("taps6", dict(taps=6)), ("taps4_nccl", dict(taps=4, table_exchange="nccl")),
                 ("taps4_4levels", dict(taps=4, force_all_levels=False))):
    if world == 1 and name.endswith("nccl"):
        continue
    try:
        out[name] = bench.extra_train_step(rank, world, dev, **kw)
    except Exception as e:  # noqa: BLE001
        out[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
if rank == 0:
    print(json.dumps(out), flush=True)
if world > 1:
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
