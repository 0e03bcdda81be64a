"""The training step of the path: forward -> losses -> backward -> gradient all-reduce -> fused AdamW.

Mirrors the arithmetic of `projects/neuralangelo/trainer.py:40-74` (loss terms and weights, coarse-to-fine /
epsilon / curvature-weight schedule) and `imaginaire/trainers/base.py:422-457` (train_step) for one model,
without the generic trainer framework (checkpointing, logging, data loading: out of scope).
"""
import torch
import torch.nn.functional as torch_F

from . import fused_chain, parallel
from .model import curvature_loss, eikonal_loss
from .optim import FusedAdamW


class TrainStep:
    def __init__(self, model, cfg=None, lr=1e-3, weight_decay=1e-2, max_iter=500000, warm_up_end=5000,
                 loss_weights=None, two_steps=(300000, 400000), gamma=10.0, use_cuda_graph=False, allow_tf32=True,
                 data_parallel=True, table_exchange="auto"):
        """data_parallel=False keeps this trainer local even inside an initialised process group (used to compute a
        single-process reference trajectory).  table_exchange: "multimem" = one fused kernel over NVLink multicast,
        "nccl" = reduce-scatter / owner AdamW / all-gather, "auto" = multimem where the fabric offers it."""
        self.model = model
        self.data_parallel = data_parallel and parallel.world_size() > 1
        self.table_exchange = table_exchange
        # imaginaire/trainers/base.py:169-178 (init_amp): TF32 library GEMMs are allowed by default.  Here that
        # only concerns the colour / background MLPs and the weight-gradient reductions; the SDF network, whose
        # finite differences need fp32-grade products, runs in the split-precision tcgen05 kernels regardless.
        torch.backends.cudnn.allow_tf32 = allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = allow_tf32
        self.max_iter, self.warm_up_end = max_iter, warm_up_end
        self.two_steps, self.gamma = two_steps, gamma
        self.base_lr = lr
        self.weights = dict(render=1.0, eikonal=0.1, curvature=5e-4)  # base.yaml:28-31
        if loss_weights:
            self.weights.update(loss_weights)
        self.init_curvature_weight = self.weights.get("curvature", 0.0)
        sdf_cfg = model.neural_sdf.cfg_sdf
        self.c2f = sdf_cfg.encoding.type == "hashgrid" and sdf_cfg.encoding.coarse2fine.enabled
        self.numerical = sdf_cfg.gradient.mode == "numerical"
        model.neural_sdf.warm_up_end = warm_up_end          # trainer.py:32
        if self.data_parallel:
            parallel.broadcast_parameters([p.data for p in model.parameters()])
        self.optim = FusedAdamW(model, lr=lr, weight_decay=weight_decay)
        self.optim.use_gathered_grads()
        self.multimem = False
        if self.data_parallel and self.optim.table is not None and self.optim.lazy and table_exchange in ("auto", "multimem"):
            self.multimem = self.optim.enable_multimem()
            if table_exchange == "multimem" and not self.multimem:
                raise RuntimeError("table_exchange='multimem' requested but the fabric offers no multicast")
        self.current_iteration = 0
        self.record_phases, self.phase_events = False, []
        self.losses, self.metrics = {}, {}
        # CUDA-graph mode: forward + loss + backward are captured once per schedule state and replayed; the
        # per-iteration scalars that torch ops consume (loss weights) live in a device tensor.
        self.use_cuda_graph = use_cuda_graph
        self._graph = self._graph_key = self._static = self._static_loss = None
        dev = next(model.parameters()).device
        self._w_dev = torch.zeros(3, dtype=torch.float32, device=dev)

    def lr_factor(self, it):
        """two_steps_with_warmup, misc.py:38-52."""
        if it < self.warm_up_end:
            return it / self.warm_up_end
        if it > self.two_steps[1]:
            return 1.0 / self.gamma ** 2
        if it > self.two_steps[0]:
            return 1.0 / self.gamma
        return 1.0

    def start_of_iteration(self, it):
        """trainer.py:63-74."""
        m = self.model
        m.set_progress(it / self.max_iter)
        sdf = m.neural_sdf
        if self.c2f:
            sdf.set_active_levels(it)
            if self.numerical:
                sdf.set_normal_epsilon()
                if "curvature" in self.weights:  # trainer.py:53-61
                    if it <= self.warm_up_end:
                        self.weights["curvature"] = it / self.warm_up_end * self.init_curvature_weight
                    else:
                        self.weights["curvature"] = self.init_curvature_weight / sdf.growth_rate ** (sdf.anneal_levels - 1)
        elif self.numerical:
            sdf.set_normal_epsilon()

    def compute_loss(self, out, image_sampled):
        """trainer.py:40-48 (train mode)."""
        # (self.losses / self.metrics keep DETACHED values: a live reference to the autograd graph would pin its
        #  AccumulateGrad nodes to the stream of this iteration and break CUDA-graph capture of a later one)
        render = torch_F.l1_loss(out["rgb"], image_sampled) * 3
        self.losses["render"] = render.detach()
        self.metrics["psnr"] = (-10 * torch_F.mse_loss(out["rgb"].detach(), image_sampled).log10())
        w = self._w_dev   # device copies of self.weights (refreshed every iteration by _upload_weights)
        total = w[0] * render
        if out["gradients"].is_cuda and "eikonal" in self.weights:
            # both regularisers in one launch per direction (csrc/glue.cu) instead of ~20 reductions / element-wise kernels
            from . import glue
            hess = out["hessians"] if "curvature" in self.weights else None
            eik, curv = glue.sdf_losses(out["gradients"], hess, out["outside"])
            self.losses["eikonal"] = eik.detach()
            total = total + w[1] * eik
            if curv is not None:
                self.losses["curvature"] = curv.detach()
                total = total + w[2] * curv
            return total
        if "eikonal" in self.weights:
            eik = eikonal_loss(out["gradients"], outside=out["outside"])
            self.losses["eikonal"] = eik.detach()
            total = total + w[1] * eik
        if "curvature" in self.weights:
            curv = curvature_loss(out["hessians"], outside=out["outside"])
            self.losses["curvature"] = curv.detach()
            total = total + w[2] * curv
        return total

    def _upload_weights(self):
        # by-value fills (kernel arguments), not an async copy from a re-used pinned buffer: the CPU may run several
        # iterations ahead of the GPU in CUDA-graph mode
        vals = (self.weights["render"], self.weights.get("eikonal", 0.0), self.weights.get("curvature", 0.0))
        if vals != getattr(self, "_w_last", None):
            for i, v in enumerate(vals):
                self._w_dev[i].fill_(float(v))
            self._w_last = vals

    def _forward_backward(self, data):
        out = self.model(data)
        loss = self.compute_loss(out, data["image_sampled"])
        loss.backward()
        self.optim.collect_grads()     # the small parameters' gradients -> flat bucket, one multi-tensor launch
        return loss

    def _schedule_key(self, data):
        """Everything that is baked into kernel arguments at capture time."""
        m = self.model
        sdf = m.neural_sdf
        return (getattr(sdf, "active_levels", None), getattr(sdf, "normal_eps", None), m.training,
                tuple((k, tuple(v.shape), v.dtype) for k, v in sorted(data.items())))

    def _graphed_forward_backward(self, data):
        key = self._schedule_key(data)
        if self._graph is None or key != self._graph_key:
            self._static = {k: v.clone() for k, v in data.items()}
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):          # warm-up outside capture (cuBLAS workspaces, lazy inits)
                for _ in range(2):
                    self._forward_backward(self._static)
            torch.cuda.current_stream().wait_stream(side)
            self.optim.zero_grad()
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph, stream=side):   # same stream as the warm-up (AccumulateGrad streams)
                self._static_loss = self._forward_backward(self._static)
            self._graph_key = key
        for k, v in data.items():
            self._static[k].copy_(v, non_blocking=True)
        self._graph.replay()
        return self._static_loss

    def step(self, data):
        """One iteration on this rank's batch: data = dict(pose, intr, idx, ray_idx, image_sampled)."""
        it = self.current_iteration
        self.start_of_iteration(it)
        self.model.train()
        self._upload_weights()
        ev = None
        if self.record_phases:     # three CUDA events per step: start / gradients ready / parameters updated
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            self.phase_events.append(ev)
            ev[0].record()
        self.optim.wait_pending()
        if self.use_cuda_graph:
            loss = self._graphed_forward_backward(data)
        else:
            loss = self._forward_backward(data)
        if ev:
            ev[1].record()
        lr = self.base_lr * self.lr_factor(it)
        world = parallel.world_size() if self.data_parallel else 1
        def repack():
            # outside any captured graph (see NeuralSDF); the colour / background operands of the chain kernel are
            # re-packed in place when training uses it, else lazily at their next graph-free use
            self.model.neural_sdf.refresh_packed_weights()
            from . import modules
            if modules.FUSED_CHAIN_IN_TRAINING:
                fused_chain.refresh_packs()
            else:
                fused_chain.invalidate_packs()

        if world > 1 and self.optim.table is not None and self.optim.lazy:
            import torch.distributed as dist
            if self.multimem:      # one kernel over NVLink multicast; small bucket + re-packing beside it
                self.optim.step_multimem(dist.get_rank(), world, lr=lr, after_small=repack)
            else:                  # reduce-scatter / owner update / all-gather
                self.optim.step_sharded(dist.get_rank(), world, lr=lr)
                repack()
        elif world == 1:
            self.optim.step(grad_scale=1.0, lr=lr, after_small=repack)
        else:
            grad_scale = parallel.allreduce_gradients(*self.optim.grad_views())
            self.optim.step(grad_scale=grad_scale, lr=lr, after_small=repack)
        self.optim.wait_pending()      # stream-level join of the side streams: whatever follows sees the whole update
        if ev:
            ev[2].record()
        self.current_iteration += 1
        return loss
