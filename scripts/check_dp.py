"""Multi-rank check (torchrun, one rank per GPU): K data-parallel training steps, every rank on its own rays, against
the SINGLE-PROCESS trajectory on the concatenated batch (DDP semantics: mean of the per-rank gradients,
imaginaire/trainers/utils/get_trainer.py:81-88).  Run for both table exchanges (fused NVLink-multicast kernel and the
NCCL reduce-scatter / all-gather form).  Rank 0 prints one JSON line per mode; keep it under profiles/.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_dp.py
"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist
from neuralangelo_b200 import parallel

rank, local, world = parallel.init_dist()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
from test_train_gpu import _batch          # noqa: E402  (tiny model + synthetic camera batch of the GPU tests)
from neuralangelo_b200.config import AttrDict
from neuralangelo_b200.model import Model
from neuralangelo_b200.trainer import TrainStep
from oracle.make_golden import small_model_config

torch.backends.cuda.matmul.allow_tf32 = False
STEPS = 14


def make(mode, data_parallel):
    torch.manual_seed(7)
    cfg = small_model_config(4)
    cfg.object.sdf.encoding.coarse2fine.step = 3          # two level activations inside the run: ownership re-cut
    model = Model(cfg, AttrDict(num_images=None, train=dict(image_size=[20, 24]), val=dict(image_size=[10, 12]))).to(dev)
    model.cfg_render.stratified = False
    ts = TrainStep(model, lr=1e-2, weight_decay=0.5, max_iter=100, warm_up_end=2, allow_tf32=False,
                   data_parallel=data_parallel, table_exchange=mode)
    return model, ts


def flat(model):
    return torch.cat([p.detach().reshape(-1).double() for p in model.parameters()])


for mode in ("multimem", "nccl"):
    try:
        model, ts = make(mode, True)
    except RuntimeError as e:      # no multicast on this fabric
        if rank == 0:
            print(json.dumps({"mode": mode, "world": world, "skipped": str(e)}))
        continue
    losses = [ts.step(_batch(dev, STEPS * 0 + world * s + rank)).item() for s in range(STEPS)]
    mine = flat(model)
    ref0 = mine.clone()
    dist.broadcast(ref0, src=0)
    same = bool(torch.equal(mine, ref0))
    # single-process reference on the concatenated batch (every rank computes it; rank 0 reports)
    ref_model, ref_ts = make("nccl", False)
    ref_losses = []
    for s in range(STEPS):
        parts = [_batch(dev, world * s + r) for r in range(world)]
        data = {k: torch.cat([p[k] for p in parts], dim=0) for k in parts[0]}
        data["idx"] = torch.arange(world, device=dev)
        ref_losses.append(ref_ts.step(data).item())
    ref = flat(ref_model)
    rel = float((mine - ref).norm() / ref.norm())
    # per-rank losses differ from the concatenated loss (own rays only); their mean over ranks must match it
    lt = torch.tensor(losses, device=dev, dtype=torch.float64)
    dist.all_reduce(lt)
    mean_losses = (lt / world).tolist()
    out = [None] * world
    dist.all_gather_object(out, (rank, same, bool(torch.isfinite(mine).all()), rel))
    if rank == 0:
        print(json.dumps({"mode": mode, "multimem_active": bool(ts.multimem), "world": world, "steps": STEPS,
                          "replicas_identical": all(o[1] for o in out), "finite": all(o[2] for o in out),
                          "rel_l2_vs_single_process": [o[3] for o in out],
                          "mean_rank_loss": [round(v, 6) for v in mean_losses],
                          "single_process_loss": [round(v, 6) for v in ref_losses],
                          "active_levels_at_end": int(model.neural_sdf.active_levels)}), flush=True)
    del model, ts, ref_model, ref_ts
    torch.cuda.empty_cache()
dist.barrier()
dist.destroy_process_group()
