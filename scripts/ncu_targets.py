"""Small workloads for ncu captures of kernels that the bench's headline does not dominate.
    python scripts/ncu_targets.py train   # 3 default-config training steps without CUDA graph (adamw, wgrad, sdf_mlp_bwd, staged renderer)
    python scripts/ncu_targets.py mesh    # inference render (chain kernel) + 256^3 extraction (lattice_sdf, mc_*)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from neuralangelo_b200 import mesh  # noqa: E402
from neuralangelo_b200.config import default_data_config, default_model_config  # noqa: E402
from neuralangelo_b200.model import Model  # noqa: E402
from neuralangelo_b200.trainer import TrainStep  # noqa: E402


def look_at(eye):
    eye = torch.tensor(eye, dtype=torch.float32)
    fwd = -eye / eye.norm()
    right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
    right = right / right.norm()
    R = torch.stack([right, torch.linalg.cross(fwd, right), fwd])
    return torch.cat([R, (-R @ eye)[:, None]], dim=1)


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "train"
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    cfg = default_model_config()
    model = Model(cfg, default_data_config()).to(dev)
    g = torch.Generator().manual_seed(0)
    intr = torch.tensor([[1111.0, 0, 400.0], [0, 1111.0, 400.0], [0, 0, 1.0]])
    if what == "train":
        ts = TrainStep(model, max_iter=500000, warm_up_end=5000, use_cuda_graph=False)
        ts.current_iteration = 5000 + 16 * 5000
        for i in range(3):
            pose = look_at((torch.randn(3, generator=g) * 3).tolist())
            ts.step(dict(pose=pose[None].to(dev), intr=intr[None].to(dev), idx=torch.tensor([i], device=dev),
                         ray_idx=torch.randint(0, 640000, (1, 512), generator=g).to(dev),
                         image_sampled=torch.rand(1, 512, 3, generator=g).to(dev)))
        torch.cuda.synchronize()
    else:
        model.eval()
        model.neural_sdf.warm_up_end = 0
        model.neural_sdf.set_active_levels(current_iter=10 ** 9)
        model.neural_sdf.set_normal_epsilon()
        with torch.no_grad():
            pose = look_at([0.0, 0.5, 3.0])
            small = torch.tensor([[80.0, 0, 32.0], [0, 80.0, 32.0], [0, 0, 1.0]])
            out = model.render_image(pose[None].to(dev), small[None].to(dev), image_size=(64, 64), stratified=False,
                                     sample_idx=torch.tensor([0], device=dev))
            m = mesh.extract_mesh(mesh.LatticeSDF(model.neural_sdf), np.array([[-1.0, 1.0]] * 3), intv=2.0 / 256, block_res=128)
        torch.cuda.synchronize()
        print("mesh", len(m.vertices), len(m.faces), None if out is None else sorted(out.keys()))


if __name__ == "__main__":
    main()
