"""CPU: pins oracle/cpu_path.py (the restatement of the reference's own CPU-runnable path, BASELINE.json
configs[0]) against the UNMODIFIED reference model run on the same weights and rays
(tests/golden/cpu_path_golden.npz from oracle/make_golden.py): outputs, loss and parameter gradients."""
import os

import numpy as np
import pytest
import torch

from oracle import cpu_path


@pytest.fixture(scope="module")
def G(golden_dir):
    z = np.load(os.path.join(golden_dir, "cpu_path_golden.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def test_cpu_path_matches_reference_model(G):
    sd = {k[3:]: v for k, v in G.items() if k.startswith("sd_")}
    path = cpu_path.CpuPath(sd, fourier_levels=6, eps=1.0 / 2048, n_coarse=16, n_fine=4, n_hier=2, requires_grad=True)
    center, ray, target = G["in_center"][0], G["in_ray"][0], G["in_target"][0]
    rands = torch.full((center.shape[0], 16), 0.5)          # stratified=False (nerf_util.py:33)
    out = path.render(center, ray, rands, progress=0.03)
    assert torch.equal(out["outside"], G["out_outside"][0])
    torch.testing.assert_close(out["dists"], G["out_dists"][0, ..., 0], rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(out["rgb"], G["out_rgb"][0], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out["weights"], G["out_weights"][0, ..., 0], rtol=1e-3, atol=1e-5)
    torch.testing.assert_close(out["gradients"], G["out_gradients"][0], rtol=1e-3, atol=2e-3)   # 1/(2 eps) = 1024
    loss = path.loss(out, target)
    assert abs(loss.item() - G["loss"].item()) <= 2e-3 * abs(G["loss"].item())
    loss.backward()
    checked = 0
    for k, p in path.w.p.items():
        ref = G.get("grad_" + k)
        if ref is None or ref.numel() == 0 or p.grad is None:
            continue
        rel = (p.grad - ref).norm().item() / max(ref.norm().item(), 1e-12)
        assert rel <= 5e-2, (k, rel)       # fp32 finite differences at eps = 1/2048 amplify rounding
        checked += 1
    assert checked >= 10


def test_cpu_path_extracts_the_geometric_init_sphere():
    """Geometric initialisation is a sphere of radius ~0.5 (mlp.py:79-84): the extracted mesh must be one."""
    path = cpu_path.CpuPath(cpu_path.geometric_init_state(hidden=64, rgb_layers=2))
    v, f, _, _ = path.extract_mesh(resolution=48, block_res=32)
    r = np.linalg.norm(v, axis=1)
    assert len(v) > 500 and 0.35 < r.min() and r.max() < 0.65 and f.max() == len(v) - 1
