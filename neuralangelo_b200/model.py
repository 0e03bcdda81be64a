"""Drop-in for `projects/neuralangelo/model.py`: same class, constructor, methods and output dictionary,
with the per-ray pipeline running on libna_b200.so kernels.

What is fused relative to the reference (reference lines in brackets):
  * hierarchical sampling: one kernel per round for alphas -> weights -> inverse-CDF samples, one rank-merge
    kernel for cat + sort + gather [model.py:232-267, ~40 ATen launches per round];
  * SDF gradients/curvature: centre + taps in one encode launch and one MLP pass [modules.py:132-175];
  * NeuS alphas ++ background alphas -> compositing weights -> RGB in one warp-per-ray kernel, with a
    hand-written backward [model.py:132-141,275-298; render.py:68-112];
  * rays are generated for the sampled pixels only [model.py:117-119].
"""
from collections import defaultdict
from functools import partial

import torch
import torch.nn.functional as torch_F

from . import camera, nerf_util, render
from .modules import BackgroundNeRF, NeuralRGB, NeuralSDF


class Model(torch.nn.Module):

    def __init__(self, cfg_model, cfg_data):
        super().__init__()
        self.cfg_render = cfg_model.render
        self.white_background = cfg_model.background.white
        self.with_background = cfg_model.background.enabled
        self.with_appear_embed = cfg_model.appear_embed.enabled
        self.anneal_end = cfg_model.object.s_var.anneal_end
        self.outside_val = 1000. * (-1 if cfg_model.object.sdf.mlp.inside_out else 1)
        self.image_size_train = cfg_data.train.image_size
        self.image_size_val = cfg_data.val.image_size
        self.progress = 1.
        # device copy of min(progress / anneal_end, 1), kept current by set_progress(): kernels read it from
        # memory, so a captured CUDA graph follows the schedule without re-capture
        self.register_buffer("_anneal_ratio", torch.ones(1, dtype=torch.float32), persistent=False)
        self.build_model(cfg_model, cfg_data)
        self.sample_dists_from_pdf = partial(nerf_util.sample_dists_from_pdf, intvs_fine=cfg_model.render.num_samples.fine)

    def build_model(self, cfg_model, cfg_data):
        if cfg_model.appear_embed.enabled:
            assert cfg_data.num_images is not None
            self.appear_embed = torch.nn.Embedding(cfg_data.num_images, cfg_model.appear_embed.dim)
            self.appear_embed_outside = torch.nn.Embedding(cfg_data.num_images, cfg_model.appear_embed.dim) \
                if cfg_model.background.enabled else None
        else:
            self.appear_embed = self.appear_embed_outside = None
        self.neural_sdf = NeuralSDF(cfg_model.object.sdf)
        self.neural_rgb = NeuralRGB(cfg_model.object.rgb, feat_dim=cfg_model.object.sdf.mlp.hidden_dim,
                                    appear_embed=cfg_model.appear_embed)
        self.background_nerf = BackgroundNeRF(cfg_model.background, appear_embed=cfg_model.appear_embed) \
            if cfg_model.background.enabled else None
        self.s_var = torch.nn.Parameter(torch.tensor(cfg_model.object.s_var.init_val, dtype=torch.float32))

    def set_progress(self, progress):
        """trainer.py:64 sets `model.progress`; this also refreshes the device-side anneal ratio (async copy)."""
        self.progress = progress
        ratio = min(progress / self.anneal_end, 1.)
        self._anneal_synced = ratio
        # fill_ passes the value as a kernel argument: asynchronous, and no host staging buffer that a CPU running
        # ahead of the GPU could overwrite before the copy has executed
        self._anneal_ratio.fill_(ratio)

    def _anneal(self):
        """The persistent device scalar when it is in sync with self.progress, else None (python value is used)."""
        expected = min(self.progress / self.anneal_end, 1.)
        if getattr(self, "_anneal_synced", None) == expected:
            return self._anneal_ratio
        return None

    def get_param_groups(self, cfg_optim):
        return self.parameters()

    def device(self):
        return next(self.parameters()).device

    # ------------------------------------------------------------------------------------------
    def forward(self, data):
        """model.py:66-71: render the pre-sampled pixels data["ray_idx"] of each image."""
        return self.render_pixels(data["pose"], data["intr"], image_size=self.image_size_train,
                                  stratified=self.cfg_render.stratified, sample_idx=data["idx"],
                                  ray_idx=data["ray_idx"])

    @torch.no_grad()
    def inference(self, data):
        """model.py:73-88."""
        self.eval()
        output = self.render_image(data["pose"], data["intr"], image_size=self.image_size_val, stratified=False,
                                   sample_idx=data["idx"])
        rot = data["pose"][..., :3, :3]
        normal_cam = -output["gradient"] @ rot.transpose(-1, -2)
        full = lambda t: t.unflatten(dim=1, sizes=tuple(self.image_size_val)).moveaxis(-1, 1)  # misc.to_full_image
        output.update(rgb_map=full(output["rgb"]), opacity_map=full(output["opacity"]), depth_map=full(output["depth"]),
                      normal_map=full(normal_cam))
        return output

    def render_image(self, pose, intr, image_size, stratified=False, sample_idx=None):
        """model.py:90-114: full image in chunks of rand_rays rays."""
        output = defaultdict(list)
        H, W = image_size
        num_rays = self.cfg_render.rand_rays
        all_idx = torch.arange(H * W, device=pose.device).repeat(pose.shape[0], 1)
        for c in range(0, H * W, num_rays):
            if pose.is_cuda and not (torch.is_grad_enabled() and (pose.requires_grad or intr.requires_grad)):
                center, ray, ray_unit, bounds = camera.ray_setup(pose, intr, image_size, all_idx[:, c:c + num_rays])
            else:
                center, ray = camera.rays_for_indices(pose, intr, image_size, all_idx[:, c:c + num_rays])
                ray_unit, bounds = torch_F.normalize(ray, dim=-1), None
            out = self.render_rays(center, ray_unit, sample_idx=sample_idx, stratified=stratified, bounds=bounds)
            if not self.training:
                dist = render.composite(out["dists"], out["weights"])
                out.update(depth=dist / ray.norm(dim=-1, keepdim=True))
            for key, value in out.items():
                if value is not None:
                    output[key].append(value.detach())
        return {key: torch.cat(value, dim=1) for key, value in output.items()}

    def render_pixels(self, pose, intr, image_size, stratified=False, sample_idx=None, ray_idx=None):
        if pose.is_cuda and not (torch.is_grad_enabled() and (pose.requires_grad or intr.requires_grad)):
            center, _, ray_unit, bounds = camera.ray_setup(pose, intr, image_size, ray_idx)   # one launch
            return self.render_rays(center, ray_unit, sample_idx=sample_idx, stratified=stratified, bounds=bounds)
        center, ray = camera.rays_for_indices(pose, intr, image_size, ray_idx)
        ray_unit = torch_F.normalize(ray, dim=-1)
        return self.render_rays(center, ray_unit, sample_idx=sample_idx, stratified=stratified)

    def render_rays(self, center, ray_unit, sample_idx=None, stratified=False, bounds=None):
        """model.py:124-156.  `bounds`: (near, far, outside) when the ray set-up kernel already produced them."""
        if bounds is not None:
            near, far, outside = bounds
        else:
            with torch.no_grad():
                near, far, outside = self.get_dist_bounds(center, ray_unit)
        app, app_outside = self.get_appearance_embedding(sample_idx, ray_unit.shape[1])
        obj = self.render_rays_object(center, ray_unit, near, far, outside, app, stratified=stratified)
        bg = self.render_rays_background(center, ray_unit, far, app_outside, stratified=stratified) \
            if self.with_background else None
        # alphas(object) ++ alphas(background) -> weights -> rgb: one kernel (model.py:132-141)
        rgb, weights, opacity_all = render.neus_render(
            ray_unit, obj["sdfs"], obj["gradients"], obj["dists"], self.s_var, obj["rgbs"], dist_far=far[..., None],
            progress=self.progress, anneal_end=self.anneal_end, anneal_ratio=self._anneal(),
            densities_bg=bg["densities"] if bg else None, dists_bg=bg["dists"] if bg else None,
            rgbs_bg=bg["rgbs"] if bg else None)
        if self.white_background:
            rgb = rgb + (1 - opacity_all)
        dists = torch.cat([obj["dists"], bg["dists"]], dim=2) if bg else obj["dists"]
        opacity = gradient = None
        if not self.training:  # model.py:169-172: object-only weights for opacity / normal
            alphas = render.neus_alphas(ray_unit, obj["sdfs"], obj["gradients"], obj["dists"], self.s_var,
                                        dist_far=far[..., None], progress=self.progress, anneal_end=self.anneal_end,
                                        anneal_ratio=self._anneal())
            w_obj = render.alpha_compositing_weights(alphas)
            opacity = render.composite(1., w_obj)
            gradient = render.composite(obj["gradients"], w_obj)
        return dict(rgb=rgb, opacity=opacity, outside=outside, dists=dists, weights=weights, gradient=gradient,
                    gradients=obj["gradients"], hessians=obj["hessians"])

    def render_rays_object(self, center, ray_unit, near, far, outside, app, stratified=False):
        """model.py:158-190 up to (not including) the alphas, which render_rays fuses with the compositing."""
        with torch.no_grad():
            dists = self.sample_dists_all(center, ray_unit, near, far, stratified=stratified)  # [B,R,N,1]
        points = camera.get_3D_points_from_dist(center, ray_unit, dists)  # [B,R,N,3]
        sdfs, feats = self.neural_sdf.forward(points)
        sdfs = torch.where(outside[..., None].expand_as(sdfs), torch.full_like(sdfs, self.outside_val), sdfs)
        gradients, hessians = self.neural_sdf.compute_gradients(points, training=self.training, sdf=sdfs)
        # normals = normalize(gradients); rgbs = neural_rgb(points, normals, rays, feats, app) (model.py:179-183)
        rgbs = self.neural_rgb.forward_from_gradients(points, gradients, ray_unit, feats, app)
        return dict(rgbs=rgbs, sdfs=sdfs, dists=dists, gradients=gradients, hessians=hessians)

    def render_rays_background(self, center, ray_unit, far, app_outside, stratified=False):
        """model.py:192-205 (densities are turned into alphas inside the fused renderer)."""
        with torch.no_grad():
            dists = self.sample_dists_background(ray_unit, far, stratified=stratified)
        points = camera.get_3D_points_from_dist(center, ray_unit, dists)
        rays_unit = ray_unit[..., None, :].expand_as(points)
        rgbs, densities = self.background_nerf.forward(points, rays_unit, app_outside)
        return dict(rgbs=rgbs, dists=dists, densities=densities)

    @torch.no_grad()
    def get_dist_bounds(self, center, ray_unit):
        """model.py:207-213: unit-sphere near/far; NaN (miss) flags `outside`, dummy range (1, 1.2)."""
        dist_near, dist_far = nerf_util.intersect_with_sphere(center, ray_unit, radius=1.)
        dist_near = dist_near.relu()
        outside = dist_near.isnan()
        dist_near = torch.where(outside, torch.ones_like(dist_near), dist_near)
        dist_far = torch.where(outside, torch.full_like(dist_far, 1.2), dist_far)
        return dist_near, dist_far, outside

    def get_appearance_embedding(self, sample_idx, num_rays):
        """model.py:215-230."""
        if not self.with_appear_embed:
            return None, None
        n_all = self.cfg_render.num_samples.coarse + self.cfg_render.num_samples.fine * self.cfg_render.num_sample_hierarchy
        app = self.appear_embed(sample_idx)[:, None, None].expand(-1, num_rays, n_all, -1)
        app_outside = None
        if self.with_background:
            app_outside = self.appear_embed_outside(sample_idx)[:, None, None].expand(
                -1, num_rays, self.cfg_render.num_samples.background, -1)
        return app, app_outside

    @torch.no_grad()
    def sample_dists_all(self, center, ray_unit, near, far, stratified=False):
        """model.py:232-248."""
        dists = nerf_util.sample_dists(ray_unit.shape[:2], dist_range=(near[..., None], far[..., None]),
                                       intvs=self.cfg_render.num_samples.coarse, stratified=stratified,
                                       device=ray_unit.device)
        n_hier = self.cfg_render.num_sample_hierarchy
        if n_hier > 0:
            sdfs = self.neural_sdf.sdf(camera.get_3D_points_from_dist(center, ray_unit, dists))
        for h in range(n_hier):
            dists_fine = self.sample_dists_hierarchical(dists, sdfs, inv_s=(64 * 2 ** h))
            if h != n_hier - 1:
                sdfs_fine = self.neural_sdf.sdf(camera.get_3D_points_from_dist(center, ray_unit, dists_fine))
                dists, sdfs = nerf_util.merge_sorted(dists, sdfs, dists_fine, sdfs_fine)
            else:
                dists, _ = nerf_util.merge_sorted(dists, None, dists_fine, None)
        return dists

    def sample_dists_hierarchical(self, dists, sdfs, inv_s, robust=True, eps=1e-5):
        """model.py:250-267."""
        assert robust and eps == 1e-5
        return nerf_util.sample_dists_hierarchical(dists, sdfs, inv_s, self.cfg_render.num_samples.fine)

    def sample_dists_background(self, ray_unit, far, stratified=False, eps=1e-5):
        """model.py:269-273."""
        inv_dists = nerf_util.sample_dists(ray_unit.shape[:2], dist_range=(1, 0),
                                           intvs=self.cfg_render.num_samples.background, stratified=stratified,
                                           device=ray_unit.device)
        return far[..., None] / (inv_dists + eps)

    def compute_neus_alphas(self, ray_unit, sdfs, gradients, dists, dist_far=None, progress=1., eps=1e-5):
        """model.py:275-292."""
        assert eps == 1e-5
        return render.neus_alphas(ray_unit, sdfs, gradients, dists, self.s_var, dist_far=dist_far, progress=progress,
                                  anneal_end=self.anneal_end)


def eikonal_loss(gradients, outside=None):
    """projects/neuralangelo/utils/misc.py:74-80."""
    err = ((gradients.norm(dim=-1) - 1.0) ** 2).nan_to_num(nan=0.0, posinf=0.0, neginf=0.0)
    return (err * (~outside).float()).mean() if outside is not None else err.mean()


def curvature_loss(hessian, outside=None):
    """projects/neuralangelo/utils/misc.py:83-89."""
    lap = hessian.sum(dim=-1).abs().nan_to_num(nan=0.0, posinf=0.0, neginf=0.0)
    return (lap * (~outside).float()).mean() if outside is not None else lap.mean()
