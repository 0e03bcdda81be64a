"""The training step of the path: forward -> losses -> backward -> gradient all-reduce -> fused AdamW.

Mirrors the arithmetic of `projects/neuralangelo/trainer.py:40-74` (loss terms and weights, coarse-to-fine /
epsilon / curvature-weight schedule) and `imaginaire/trainers/base.py:422-457` (train_step) for one model,
without the generic trainer framework (checkpointing, logging, data loading: out of scope).
"""
import torch
import torch.nn.functional as torch_F

from . import parallel
from .model import curvature_loss, eikonal_loss
from .optim import FusedAdamW


class TrainStep:
    def __init__(self, model, cfg=None, lr=1e-3, weight_decay=1e-2, max_iter=500000, warm_up_end=5000,
                 loss_weights=None, two_steps=(300000, 400000), gamma=10.0):
        self.model = model
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
        parallel.broadcast_parameters([p.data for p in model.parameters()])
        self.optim = FusedAdamW(model, lr=lr, weight_decay=weight_decay)
        self.current_iteration = 0
        self.losses, self.metrics = {}, {}

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
        m.progress = it / self.max_iter
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
        self.losses["render"] = torch_F.l1_loss(out["rgb"], image_sampled) * 3
        self.metrics["psnr"] = -10 * torch_F.mse_loss(out["rgb"], image_sampled).log10()
        total = self.weights["render"] * self.losses["render"]
        if "eikonal" in self.weights:
            self.losses["eikonal"] = eikonal_loss(out["gradients"], outside=out["outside"])
            total = total + self.weights["eikonal"] * self.losses["eikonal"]
        if "curvature" in self.weights:
            self.losses["curvature"] = curvature_loss(out["hessians"], outside=out["outside"])
            total = total + self.weights["curvature"] * self.losses["curvature"]
        return total

    def step(self, data):
        """One iteration on this rank's batch: data = dict(pose, intr, idx, ray_idx, image_sampled)."""
        it = self.current_iteration
        self.start_of_iteration(it)
        self.model.train()
        out = self.model(data)
        loss = self.compute_loss(out, data["image_sampled"])
        loss.backward()
        grad_scale = parallel.allreduce_gradients(*self.optim.grad_views())
        self.optim.step(grad_scale=grad_scale, lr=self.base_lr * self.lr_factor(it))
        self.current_iteration += 1
        return loss.detach()
