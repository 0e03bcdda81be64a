"""Generates tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference
through oracle/reference_shim.py) on seeded inputs.  Run in the authoring container only:

    python -m oracle.make_golden

The fixtures are small (a few hundred KB) and committed; tests/test_oracle_render.py checks
oracle/render_ref.py against them, so the oracle stays pinned on machines without the reference.
"""
import os
import types

import numpy as np
import torch

from . import reference_shim

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy()


def render_fixtures():
    reference_shim.install()
    from projects.nerf.utils import render, nerf_util
    from projects.neuralangelo.model import Model

    g = torch.Generator().manual_seed(20231106)
    B, R, N, NB, NF = 2, 24, 37, 9, 16
    out = {}
    # ---- render.py functions ------------------------------------------------------------
    alphas = torch.rand(B, R, N, generator=g) ** 3
    alphas[0, 0, 5] = 1.0   # exact-one alpha: transmittance hits 0 (clip at model.py:290 makes this reachable)
    alphas[0, 1, :] = 0.0
    out["aw_alphas"] = _np(alphas)
    out["aw_weights"] = _np(render.alpha_compositing_weights(alphas))
    q = torch.randn(B, R, N, 3, generator=g)
    out["comp_q"] = _np(q)
    out["comp_out"] = _np(render.composite(q, render.alpha_compositing_weights(alphas)))
    dens = torch.rand(B, R, NB, generator=g) * 5
    dists = torch.sort(torch.rand(B, R, NB, 1, generator=g) * 4 + 0.5, dim=2).values
    far = dists[..., -1:, :] + torch.rand(B, R, 1, 1, generator=g)
    out["da_dens"], out["da_dists"], out["da_far"] = _np(dens), _np(dists), _np(far)
    out["da_alphas_far"] = _np(render.volume_rendering_alphas_dist(dens, dists, far))
    out["da_alphas_nofar"] = _np(render.volume_rendering_alphas_dist(dens, dists, None))
    # ---- nerf_util.sample_dists_from_pdf -------------------------------------------------
    bins = torch.sort(torch.rand(B, R, N, 1, generator=g) * 3 + 1, dim=2).values
    w = torch.rand(B, R, N - 1, generator=g) ** 4
    w[0, 2, 10:20] = 0.0    # flat CDF stretch
    w[1, 3, :] = 0.0        # all-zero weights
    out["pdf_bins"], out["pdf_w"] = _np(bins), _np(w)
    out["pdf_out"] = _np(nerf_util.sample_dists_from_pdf(bins, w, NF))
    # ---- Model methods (unbound: they only touch self.s_var / self.anneal_end / self.sample_dists_from_pdf)
    stub = types.SimpleNamespace()
    stub.s_var = torch.tensor(3.0)
    stub.anneal_end = 0.1
    stub._get_iter_cos = lambda c, progress=1.: Model._get_iter_cos(stub, c, progress=progress)
    import functools
    stub.sample_dists_from_pdf = functools.partial(nerf_util.sample_dists_from_pdf, intvs_fine=NF)
    ray = torch.nn.functional.normalize(torch.randn(B, R, 3, generator=g), dim=-1)
    sd = torch.sort(torch.rand(B, R, N, 1, generator=g) * 2 + 1, dim=2).values
    # SDF crossing zero along the ray plus noise; gradients roughly anti-parallel to the ray
    sdfs = (2.0 - sd) * 0.4 + 0.02 * torch.randn(B, R, N, 1, generator=g)
    grads = -ray[:, :, None, :] * (0.8 + 0.4 * torch.rand(B, R, N, 1, generator=g)) + \
        0.3 * torch.randn(B, R, N, 3, generator=g)
    dfar = sd[..., -1:, :] + 0.05 + torch.rand(B, R, 1, 1, generator=g) * 0.1
    out["na_ray"], out["na_sdf"], out["na_grad"], out["na_dists"], out["na_far"] = map(_np, (ray, sdfs, grads, sd, dfar))
    for tag, prog in (("p0", 0.0), ("p05", 0.05), ("p1", 1.0)):
        out[f"na_alphas_{tag}"] = _np(Model.compute_neus_alphas(stub, ray, sdfs, grads, sd, dist_far=dfar, progress=prog))
    out["na_alphas_nofar"] = _np(Model.compute_neus_alphas(stub, ray, sdfs, grads, sd, dist_far=None, progress=1.0))
    # gradients of a scalar loss through compute_neus_alphas -> weights -> composite
    sdfs_g, grads_g = sdfs.clone().requires_grad_(True), grads.clone().requires_grad_(True)
    stub.s_var = torch.tensor(3.0, requires_grad=True)
    rgbs = torch.rand(B, R, N, 3, generator=g)
    a = Model.compute_neus_alphas(stub, ray, sdfs_g, grads_g, sd, dist_far=dfar, progress=0.05)
    wts = render.alpha_compositing_weights(a)
    rgb = render.composite(rgbs, wts)
    tgt = torch.rand(B, R, 3, generator=g)
    loss = ((rgb - tgt) ** 2).sum()
    loss.backward()
    out["bw_rgbs"], out["bw_tgt"], out["bw_rgb"] = _np(rgbs), _np(tgt), _np(rgb)
    out["bw_d_sdf"], out["bw_d_grad"], out["bw_d_svar"] = _np(sdfs_g.grad), _np(grads_g.grad), _np(stub.s_var.grad)
    stub.s_var = torch.tensor(3.0)
    # hierarchical upsampling
    for h in (0, 2):
        out[f"up_fine_h{h}"] = _np(Model.sample_dists_hierarchical(stub, sd, sdfs, inv_s=64 * 2 ** h))
    # known answers of SURVEY.md section 4
    kat_s = torch.tensor([.3, .1, -.1, -.3]).view(1, 1, 4, 1)
    kat_d = torch.tensor([1, 1.2, 1.4, 1.6]).view(1, 1, 4, 1)
    NF4 = functools.partial(nerf_util.sample_dists_from_pdf, intvs_fine=4)
    stub.sample_dists_from_pdf = NF4
    out["kat_up"] = _np(Model.sample_dists_hierarchical(stub, kat_d, kat_s, inv_s=64))
    np.savez_compressed(os.path.join(OUT, "render_golden.npz"), **out)
    print("wrote render_golden.npz with", len(out), "arrays")


def small_model_config(taps):
    """A scaled-down base.yaml `model:` section (4 hash-grid levels, 64-wide MLPs) so that the fixture stays small."""
    from neuralangelo_b200.config import AttrDict
    return AttrDict(
        object=dict(
            sdf=dict(mlp=dict(num_layers=1, hidden_dim=64, skip=[], activ="softplus", activ_params=dict(beta=100),
                              geometric_init=True, weight_norm=True, out_bias=0.5, inside_out=False),
                     encoding=dict(type="hashgrid", levels=4,
                                   hashgrid=dict(min_logres=2, max_logres=5, dict_size=10, dim=4, range=[-2, 2]),
                                   coarse2fine=dict(enabled=True, init_active_level=2, step=10)),
                     gradient=dict(mode="numerical", taps=taps)),
            rgb=dict(mlp=dict(num_layers=2, hidden_dim=64, skip=[], activ="relu_", activ_params={}, weight_norm=True),
                     mode="idr", encoding_view=dict(type="spherical", levels=3)),
            s_var=dict(init_val=3.0, anneal_end=0.1)),
        background=dict(enabled=True, white=False,
                        mlp=dict(num_layers=4, hidden_dim=32, skip=[2], num_layers_rgb=2, hidden_dim_rgb=16, skip_rgb=[],
                                 activ="relu", activ_params={}, activ_density="softplus", activ_density_params={}),
                        view_dep=True, encoding=dict(type="fourier", levels=4),
                        encoding_view=dict(type="spherical", levels=3)),
        render=dict(rand_rays=24, num_samples=dict(coarse=8, fine=4, background=6), num_sample_hierarchy=2,
                    stratified=False),
        appear_embed=dict(enabled=False, dim=8))


def model_fixtures():
    """Full `Model.forward` (train mode) + loss gradients of the UNMODIFIED reference model on the CPU, hash grid
    provided by oracle/tcnn_stub.py.  Stores inputs, the complete state dict, outputs and parameter gradients."""
    from . import tcnn_stub
    reference_shim.install(tcnn_module=tcnn_stub)
    from neuralangelo_b200.config import AttrDict
    from projects.neuralangelo.model import Model
    from projects.neuralangelo.utils.misc import eikonal_loss, curvature_loss

    for taps in (4, 6):
        torch.manual_seed(1234 + taps)
        cfg_model = small_model_config(taps)
        cfg_data = AttrDict(num_images=None, train=dict(image_size=[20, 24]), val=dict(image_size=[10, 12]))
        model = Model(cfg_model, cfg_data)
        # make the encoding matter: geometric init zeroes its first-layer weights and tcnn's init is +-1e-4
        with torch.no_grad():
            model.neural_sdf.tcnn_encoding.params.uniform_(-0.3, 0.3)
            lin0 = model.neural_sdf.mlp.linears[0]
            lin0.weight_v[:, 3:] = torch.randn_like(lin0.weight_v[:, 3:]) * 0.05
        model.train()
        model.neural_sdf.warm_up_end = 5
        model.neural_sdf.set_active_levels(35)      # -> anneal 3, active 3 of 4 levels
        model.neural_sdf.set_normal_epsilon()
        model.progress = 0.04
        # two cameras on a radius-3 sphere looking at the origin, wide enough that some rays miss the unit sphere
        def look_at(eye):
            eye = torch.tensor(eye, dtype=torch.float32)
            fwd = -eye / eye.norm()
            right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
            right = right / right.norm()
            up = torch.linalg.cross(fwd, right)
            R = torch.stack([right, up, fwd], dim=0)    # world -> camera rotation
            return torch.cat([R, (-R @ eye)[:, None]], dim=1)
        pose = torch.stack([look_at([0.3, 0.2, -3.0]), look_at([2.0, 0.5, 2.2])])
        intr = torch.tensor([[36.0, 0.0, 12.0], [0.0, 36.0, 10.0], [0.0, 0.0, 1.0]]).repeat(2, 1, 1)
        g = torch.Generator().manual_seed(99)
        ray_idx = torch.stack([torch.randperm(20 * 24, generator=g)[:24] for _ in range(2)])
        target = torch.rand(2, 24, 3, generator=g)
        if taps == 4:
            # whole forward() incl. ray generation, one image (with B > 1 the reference's own `.view(-1, 3)` at
            # modules.py:84 fails on the CPU: slice_by_ray_idx returns a ray-major layout there)
            pose, intr, ray_idx, target = pose[:1], intr[:1], ray_idx[:1], target[:1]
            out = model(dict(pose=pose, intr=intr, idx=torch.arange(1), ray_idx=ray_idx))
        else:
            # two images through render_rays with explicitly contiguous rays
            from projects.nerf.utils import camera as ref_camera, nerf_util as ref_nerf_util
            center, ray = ref_camera.get_center_and_ray(pose, intr, [20, 24])
            center = ref_nerf_util.slice_by_ray_idx(center, ray_idx).contiguous()
            ray = ref_nerf_util.slice_by_ray_idx(ray, ray_idx).contiguous()
            out = model.render_rays(center, torch.nn.functional.normalize(ray, dim=-1), sample_idx=torch.arange(2),
                                    stratified=False)
        loss = 3 * torch.nn.functional.l1_loss(out["rgb"], target) + 0.1 * eikonal_loss(out["gradients"], out["outside"]) \
            + 5e-4 * curvature_loss(out["hessians"], out["outside"])
        loss.backward()
        fx = {"in_pose": _np(pose), "in_intr": _np(intr), "in_ray_idx": _np(ray_idx), "in_target": _np(target),
              "loss": _np(loss), "normal_eps": np.float64(model.neural_sdf.normal_eps),
              "active_levels": np.int64(model.neural_sdf.active_levels)}
        for k in ("rgb", "dists", "weights", "gradients", "hessians", "outside"):
            fx["out_" + k] = _np(out[k])
        for k, v in model.state_dict().items():
            fx["sd_" + k] = _np(v)
        for k, p in model.named_parameters():
            fx["grad_" + k] = _np(p.grad) if p.grad is not None else np.zeros(0, dtype=np.float32)
        np.savez_compressed(os.path.join(OUT, f"model_golden_taps{taps}.npz"), **fx)
        print(f"wrote model_golden_taps{taps}.npz: loss {float(loss):.6f}, outside rays {int(out['outside'].sum())},",
              "table grad nnz", int((model.neural_sdf.tcnn_encoding.params.grad != 0).sum()))


def _look_at(eye):
    eye = torch.tensor(eye, dtype=torch.float32)
    fwd = -eye / eye.norm()
    right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
    right = right / right.norm()
    up = torch.linalg.cross(fwd, right)
    R = torch.stack([right, up, fwd], dim=0)    # world -> camera rotation
    return torch.cat([R, (-R @ eye)[:, None]], dim=1)


DEFAULT_CASES = ((4, 16), (6, 16), (4, 9))   # (taps, active levels) of the default-config fixtures


def default_model_fixtures(cases=DEFAULT_CASES, n_rays=16):
    """The UNMODIFIED reference model at base.yaml's DEFAULT shape (16 levels x 8 features, dict_size 22, 256-wide
    MLPs, 64 + 4 x 16 object and 32 background samples) on the CPU, hash grid by oracle/tcnn_stub.py: train-mode
    render_rays + loss + backward.  Too large to store whole -- see oracle/golden_recipe.py: the table and the large
    weight tensors are recorded as recipes; the table gradient as per-level checksums plus a sample of entries; large
    MLP gradients as norm + a sample."""
    from . import tcnn_stub, golden_recipe as gr
    reference_shim.install(tcnn_module=tcnn_stub)
    from neuralangelo_b200.config import AttrDict, default_model_config
    from projects.neuralangelo.model import Model
    from projects.neuralangelo.utils.misc import eikonal_loss, curvature_loss

    for taps, active in cases:
        torch.manual_seed(4321 + taps)
        cfg_model = default_model_config()
        cfg_model.object.sdf.gradient.taps = taps
        cfg_model.render.rand_rays = n_rays
        cfg_model.render.stratified = False
        cfg_data = AttrDict(num_images=None, train=dict(image_size=[40, 48]), val=dict(image_size=[10, 12]))
        model = Model(cfg_model, cfg_data)
        fx = {}
        # ---- parameters: big tensors from recipes (same statistics as the reference's own initialisation) ----
        table_seed, table_amp = 777 + taps, 0.05
        enc = model.neural_sdf.tcnn_encoding
        with torch.no_grad():
            enc.params.copy_(torch.from_numpy(gr.table_tensor(table_seed, enc.params.numel(), table_amp)))
            seed = 1000 * taps + active
            for name, p in sorted(model.named_parameters()):
                if p is enc.params or p.numel() < gr.BIG:
                    continue
                seed += 1
                mean, std = float(p.mean()), float(p.std())
                if name == "neural_sdf.mlp.linears.0.weight_v":
                    # geometric init zeroes the encoding columns; give them weight so that the hash grid matters
                    w = gr.normal_tensor(seed, p.shape, 0.0, 1.0)
                    xyz_std = float(p[:, :3].std())
                    w[:, :3] *= xyz_std
                    w[:, 3:] *= 0.02
                    p.copy_(torch.from_numpy(w))
                    fx["recipe_" + name] = np.array([seed, -1.0, xyz_std, 0.02], dtype=np.float64)
                else:
                    p.copy_(torch.from_numpy(gr.normal_tensor(seed, p.shape, mean, std)))
                    fx["recipe_" + name] = np.array([seed, mean, std, 0.0], dtype=np.float64)
        fx["table_recipe"] = np.array([table_seed, table_amp, enc.params.numel()], dtype=np.float64)
        for k, v in model.state_dict().items():
            if v.numel() < gr.BIG:
                fx["sd_" + k] = _np(v)
        model.train()
        sdf = model.neural_sdf
        sdf.warm_up_end = 5000
        it = 5000 + active * 5000            # anneal = active (modules.py:97-100)
        sdf.set_active_levels(it)
        sdf.set_normal_epsilon()
        assert sdf.active_levels == active
        model.progress = 0.04
        pose = torch.stack([_look_at([0.3, 0.2, -3.0])])
        intr = torch.tensor([[70.0, 0.0, 24.0], [0.0, 70.0, 20.0], [0.0, 0.0, 1.0]]).repeat(1, 1, 1)
        g = torch.Generator().manual_seed(99 + taps)
        ray_idx = torch.stack([torch.randperm(40 * 48, generator=g)[:n_rays]])
        target = torch.rand(1, n_rays, 3, generator=g)
        out = model(dict(pose=pose, intr=intr, idx=torch.arange(1), ray_idx=ray_idx))
        loss = 3 * torch.nn.functional.l1_loss(out["rgb"], target) + 0.1 * eikonal_loss(out["gradients"], out["outside"]) \
            + 5e-4 * curvature_loss(out["hessians"], out["outside"])
        loss.backward()
        fx.update({"in_pose": _np(pose), "in_intr": _np(intr), "in_ray_idx": _np(ray_idx), "in_target": _np(target),
                   "loss": _np(loss), "normal_eps": np.float64(sdf.normal_eps), "active_levels": np.int64(active),
                   "iteration": np.int64(it), "taps": np.int64(taps)})
        for k in ("rgb", "dists", "weights", "gradients", "hessians", "outside"):
            fx["out_" + k] = _np(out[k])
        # ---- reference gradients (kept before the sensitivity run below re-uses .grad) ----
        ref_grads = {name: (p.grad.detach().clone() if p.grad is not None else None) for name, p in model.named_parameters()}
        # ---- sensitivity of the reference's own outputs and gradients to the SDF tolerance ----
        # The fused SDF network is specified to 2e-6 absolute (tests/test_mlp_gpu.py).  At the default shape the
        # finite-difference step is 1/2049, so that error is amplified into the gradients (x 1/eps) and from there
        # into alphas / weights / loss gradients.  Measure it on the reference itself: same sample positions, every
        # SDF value perturbed by U(-2e-6, 2e-6) (additive: the autograd graph is unchanged); the GPU test allows its
        # base tolerance + 4 x this movement.
        n_obj = out["dists"].shape[2] - cfg_model.render.num_samples.background
        fixed = out["dists"][:, :, :n_obj].detach().clone()
        orig_sampler, orig_mlp = model.sample_dists_all, sdf.mlp.forward
        gen = torch.Generator().manual_seed(5)

        def noisy(*a, **kw):
            s_, f_ = orig_mlp(*a, **kw)
            if s_ is not None:
                s_ = s_ + (torch.rand(s_.shape, generator=gen) * 2 - 1) * 2e-6
            return s_, f_
        model.sample_dists_all = lambda *a, **k: fixed
        sdf.mlp.forward = noisy
        model.zero_grad(set_to_none=True)
        pert = model(dict(pose=pose, intr=intr, idx=torch.arange(1), ray_idx=ray_idx))
        loss_p = 3 * torch.nn.functional.l1_loss(pert["rgb"], target) \
            + 0.1 * eikonal_loss(pert["gradients"], pert["outside"]) + 5e-4 * curvature_loss(pert["hessians"], pert["outside"])
        loss_p.backward()
        model.sample_dists_all, sdf.mlp.forward = orig_sampler, orig_mlp
        inside = ~out["outside"][..., 0]
        for k in ("rgb", "weights", "gradients", "hessians"):
            fx["sens_" + k] = np.float64((pert[k][inside] - out[k][inside]).detach().abs().max())
        fx["sens_loss"] = np.float64(abs(float(loss_p) - float(loss)))
        F_ = enc.tbl["n_features"]
        for name, p in model.named_parameters():
            g0, g1 = ref_grads[name], p.grad
            if g0 is None or g1 is None:
                continue
            if p is enc.params:
                rel = []
                for lv in enc.tbl["levels"]:
                    sl_ = slice(lv["offset"] * F_, (lv["offset"] + lv["entries"]) * F_)
                    rel.append(float((g1[sl_] - g0[sl_]).double().norm() / max(float(g0[sl_].double().norm()), 1e-30)))
                fx["tg_sens_rel"] = np.array(rel)
            else:
                fx["gsens_" + name] = np.float64((g1 - g0).double().norm() / max(float(g0.double().norm()), 1e-30))
        for name, p in model.named_parameters():   # restore the reference gradients for the records below
            p.grad = ref_grads[name]
        # ---- gradients ----
        for name, p in model.named_parameters():
            if p is enc.params:
                continue
            gflat = p.grad.detach().flatten() if p.grad is not None else torch.zeros(0)
            if p.numel() < gr.BIG:
                fx["grad_" + name] = _np(gflat.view_as(p)) if p.grad is not None else np.zeros(0, dtype=np.float32)
            else:
                pos = gr.sample_positions(p.numel(), 4096, seed=17)
                fx["gradnorm_" + name] = np.float64(gflat.double().norm())
                fx["gradsample_" + name] = _np(gflat[torch.from_numpy(pos)])
        tg = enc.params.grad.detach()
        F = enc.tbl["n_features"]
        lv_sum, lv_abs, lv_l2, lv_nnz = [], [], [], []
        for lv in enc.tbl["levels"]:
            part = tg[lv["offset"] * F:(lv["offset"] + lv["entries"]) * F].double()
            lv_sum.append(float(part.sum())), lv_abs.append(float(part.abs().sum()))
            lv_l2.append(float(part.norm())), lv_nnz.append(int((part != 0).sum()))
        fx["tg_level_sum"], fx["tg_level_abs"] = np.array(lv_sum), np.array(lv_abs)
        fx["tg_level_l2"], fx["tg_level_nnz"] = np.array(lv_l2), np.array(lv_nnz, dtype=np.int64)
        nz = torch.nonzero(tg).flatten()
        pick = nz[torch.from_numpy(gr.sample_positions(nz.numel(), 20000, seed=23))]
        fx["tg_index"], fx["tg_value"] = _np(pick), _np(tg[pick])
        # entries that must stay exactly zero (sample of untouched positions)
        zpos = torch.from_numpy(gr.sample_positions(tg.numel(), 20000, seed=29))
        zpos = zpos[tg[zpos] == 0]
        fx["tg_zero_index"] = _np(zpos)
        name = f"default_model_golden_taps{taps}_active{active}.npz"
        np.savez_compressed(os.path.join(OUT, name), **fx)
        print(f"wrote {name}: loss {float(loss):.6f}, outside rays {int(out['outside'].sum())}, table grad nnz "
              f"{int(nz.numel())}, size {os.path.getsize(os.path.join(OUT, name)) / 1e3:.0f} kB")


TEX_GAIN = 12.0   # factor on the colour head's weight_g for the texture fixture (applied on both sides)


def inference_texture_fixtures():
    """`Model.inference` (model.py:73-88: rgb / opacity / depth / normal maps of a validation view) and
    `extract_texture` (utils/mesh.py:52-65) of the UNMODIFIED reference on the small golden models, weights taken from
    model_golden_taps{4,6}.npz.  extract_texture hard-codes `.cuda()`; it is run with Tensor.cuda patched to the
    identity, nothing else touched."""
    from . import tcnn_stub
    reference_shim.install(tcnn_module=tcnn_stub)
    from neuralangelo_b200.config import AttrDict
    from projects.neuralangelo.model import Model
    from projects.neuralangelo.utils import mesh as ref_mesh
    fx = {}
    for taps in (4, 6):
        z = np.load(os.path.join(OUT, f"model_golden_taps{taps}.npz"))
        cfg_data = AttrDict(num_images=None, train=dict(image_size=[20, 24]), val=dict(image_size=[10, 12]))
        model = Model(small_model_config(taps), cfg_data)
        model.load_state_dict({k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd_")}, strict=True)
        model.neural_sdf.warm_up_end = 5
        model.neural_sdf.set_active_levels(35)
        model.neural_sdf.set_normal_epsilon()
        model.progress = 0.04
        pose, intr = torch.from_numpy(z["in_pose"])[:1], torch.from_numpy(z["in_intr"])[:1]
        intr_val = torch.tensor([[0.5, 0, 0], [0, 0.5, 0], [0, 0, 1.0]]) @ intr
        out = model.inference(dict(pose=pose, intr=intr_val, idx=torch.arange(1)))
        for k in ("rgb_map", "opacity_map", "depth_map", "normal_map"):
            fx[f"inf{taps}_{k}"] = _np(out[k])
        fx[f"inf{taps}_pose"], fx[f"inf{taps}_intr"] = _np(pose), _np(intr_val)
        # texture: points near the model's zero level set (|x| ~ 0.5 after geometric init) and a few far from it
        g = torch.Generator().manual_seed(50 + taps)
        dirs = torch.nn.functional.normalize(torch.randn(160, 3, generator=g), dim=-1)
        xyz = (dirs * (0.35 + 0.3 * torch.rand(160, 1, generator=g))).numpy().astype(np.float64)
        with torch.no_grad():   # spread the colours: the random-init colour head outputs ~0.5 everywhere
            model.neural_rgb.mlp.linears[-1].weight_g.mul_(TEX_GAIN)
        orig_cuda = torch.Tensor.cuda
        torch.Tensor.cuda = lambda self, *a, **k: self
        try:
            colors = ref_mesh.extract_texture(xyz, model.neural_rgb, model.neural_sdf, None)
        finally:
            torch.Tensor.cuda = orig_cuda
        fx[f"tex{taps}_xyz"], fx[f"tex{taps}_colors"] = xyz, colors
        print(f"taps {taps}: opacity range {float(out['opacity_map'].min()):.3f}..{float(out['opacity_map'].max()):.3f}, "
              f"colour range {colors.min()}..{colors.max()}")
    np.savez_compressed(os.path.join(OUT, "inference_texture_golden.npz"), **fx)
    print("wrote inference_texture_golden.npz")


def cpu_path_fixtures():
    """BASELINE.json configs[0] shape (Fourier-encoded NeuralSDF, no background, 6-tap gradients) on the UNMODIFIED
    reference model, 64-wide MLPs to keep the fixture small: state dict, rays, outputs, loss, parameter gradients."""
    reference_shim.install()
    from neuralangelo_b200.config import AttrDict
    from projects.neuralangelo.model import Model
    from projects.neuralangelo.utils.misc import eikonal_loss, curvature_loss
    torch.manual_seed(7)
    cfg = small_model_config(6)
    cfg.object.sdf.encoding = AttrDict(type="fourier", levels=6, coarse2fine=dict(enabled=False, init_active_level=4, step=5000))
    cfg.background.enabled = False
    cfg.render.num_samples = AttrDict(coarse=16, fine=4, background=0)
    model = Model(cfg, AttrDict(num_images=None, train=dict(image_size=[20, 24]), val=dict(image_size=[10, 12])))
    with torch.no_grad():   # give the positional-encoding columns some weight so that they matter
        lin0 = model.neural_sdf.mlp.linears[0]
        lin0.weight_v[:, 3:] = torch.randn_like(lin0.weight_v[:, 3:]) * 0.02
    model.train()
    model.neural_sdf.normal_eps = 1.0 / 2048      # set by hand: set_normal_epsilon needs hash-grid resolutions (modules.py:102-107)
    model.progress = 0.03
    g = torch.Generator().manual_seed(5)
    R = 40
    d = torch.randn(1, R, 3, generator=g) * 0.15 + torch.tensor([0.0, 0.0, 1.0])
    d[0, :3] = torch.tensor([0.9, 0.1, 0.3])      # rays that miss the unit sphere
    ray = torch.nn.functional.normalize(d, dim=-1)
    center = torch.tensor([0.0, 0.0, -3.0]).expand(1, R, 3).contiguous()
    target = torch.rand(1, R, 3, generator=g)
    out = model.render_rays(center, ray, sample_idx=torch.arange(1), stratified=False)
    loss = 3 * torch.nn.functional.l1_loss(out["rgb"], target) + 0.1 * eikonal_loss(out["gradients"], out["outside"]) \
        + 5e-4 * curvature_loss(out["hessians"], out["outside"])
    loss.backward()
    fx = {"in_center": _np(center), "in_ray": _np(ray), "in_target": _np(target), "loss": _np(loss)}
    for k in ("rgb", "dists", "weights", "gradients", "hessians", "outside"):
        fx["out_" + k] = _np(out[k])
    for k, v in model.state_dict().items():
        fx["sd_" + k] = _np(v)
    for k, p in model.named_parameters():
        fx["grad_" + k] = _np(p.grad) if p.grad is not None else np.zeros(0, dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, "cpu_path_golden.npz"), **fx)
    print(f"wrote cpu_path_golden.npz: loss {float(loss):.6f}, outside rays {int(out['outside'].sum())}")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    import sys
    if "--default-only" not in sys.argv:
        render_fixtures()
        model_fixtures()
        cpu_path_fixtures()
        inference_texture_fixtures()
    if "--default" in sys.argv or "--default-only" in sys.argv:   # slow (1.46 GB table, dense autograd on the CPU)
        default_model_fixtures()
