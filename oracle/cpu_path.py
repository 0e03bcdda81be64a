"""CPU restatement of the reference's own CPU-runnable path (BASELINE.json configs[0]) -- TEST INFRASTRUCTURE and
the timed CPU baseline of the rendering / extraction path.

Fourier-encoded NeuralSDF (projects/neuralangelo/utils/modules.py:68-95,132-156 with encoding.type=fourier),
IDR colour MLP (modules.py:209-229), NeuS sampling and rendering (projects/neuralangelo/model.py:124-298) without
background, and blocked extraction with the C marching cubes (projects/neuralangelo/utils/mesh.py:24-133).
Pinned: tests/test_oracle_cpu_path.py compares it with outputs of the UNMODIFIED reference model run on the same
weights (tests/golden/cpu_path_golden.npz, produced by oracle/make_golden.py).  Plain torch on the host; weights are
read from a reference-format state dict (keys neural_sdf.mlp.linears.{i}.{weight_g,weight_v,bias}, ...).
"""
import math
import time

import numpy as np
import torch
import torch.nn.functional as F

from . import mcubes_ref
from . import render_ref as rr

SH_C0, SH_C1 = 0.28209479177387814, 0.4886025119029199
SH_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
SH_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
         1.445305721320277, -0.5900435899266435)


def fourier(x, levels):
    """nerf_util.py:134-147: per input dim [sin(2^l pi x) for l] then [cos ...]."""
    freq = 2 ** torch.arange(levels, dtype=torch.float32) * math.pi
    spec = x[..., None] * freq
    return torch.stack([spec.sin(), spec.cos()], dim=-2).reshape(*x.shape[:-1], -1)


def sh3(d):
    """spherical_harmonics.py:47-72, degree 3 (16 values)."""
    x, y, z = d.unbind(-1)
    xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
    return torch.stack([torch.full_like(x, SH_C0), -SH_C1 * y, SH_C1 * z, -SH_C1 * x, SH_C2[0] * xy, SH_C2[1] * yz,
                        SH_C2[2] * (2 * zz - xx - yy), SH_C2[3] * xz, SH_C2[4] * (xx - yy), SH_C3[0] * y * (3 * xx - yy),
                        SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy), SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy),
                        SH_C3[4] * x * (4 * zz - xx - yy), SH_C3[5] * z * (xx - yy), SH_C3[6] * x * (xx - 3 * yy)], dim=-1)


class Weights:
    """Leaf tensors of a reference state dict, with weight norm applied on the fly (mlp.py:42-43)."""

    def __init__(self, state_dict, requires_grad=False):
        self.p = {k: v.clone().float().requires_grad_(requires_grad) for k, v in state_dict.items()}

    def linear(self, prefix, x):
        if prefix + ".weight_v" in self.p:
            v, g = self.p[prefix + ".weight_v"], self.p[prefix + ".weight_g"]
            w = v * (g / v.norm(dim=1, keepdim=True))
        else:
            w = self.p[prefix + ".weight"]
        return F.linear(x, w, self.p[prefix + ".bias"])

    def count(self, prefix):
        return len({k[len(prefix):].split(".")[1] for k in self.p if k.startswith(prefix + ".")})


class CpuPath:
    def __init__(self, state_dict, fourier_levels=6, eps=1.0 / 2048, n_coarse=64, n_fine=16, n_hier=2, anneal_end=0.1,
                 requires_grad=False):
        self.w = Weights(state_dict, requires_grad)
        self.levels, self.eps = fourier_levels, eps
        self.n_coarse, self.n_fine, self.n_hier, self.anneal_end = n_coarse, n_fine, n_hier, anneal_end

    # mlp.py:55-69 with num_layers = 1: sdf head reads the first hidden layer
    def sdf_feat(self, pts, with_feat=True):
        x = torch.cat([pts, fourier(pts, self.levels)], dim=-1)
        h = F.softplus(self.w.linear("neural_sdf.mlp.linears.0", x), beta=100)
        sdf = self.w.linear("neural_sdf.mlp.linear_sdf", h)
        feat = F.softplus(self.w.linear("neural_sdf.mlp.linears.1", h), beta=100) if with_feat else None
        return sdf, feat

    def gradients(self, pts, sdf):
        """6-tap central differences + diagonal hessian, modules.py:132-156."""
        e = self.eps
        g, hs = [], []
        for axis in range(3):
            off = torch.zeros(3)
            off[axis] = e
            sp, sm = self.sdf_feat(pts + off, False)[0], self.sdf_feat(pts - off, False)[0]
            g.append((sp - sm) / (2 * e))
            hs.append((sp + sm - 2 * sdf) / (e ** 2))
        return torch.cat(g, dim=-1), torch.cat(hs, dim=-1)

    def rgb(self, pts, normals, rays, feats):
        x = torch.cat([pts, sh3(rays), normals, feats], dim=-1)
        n = self.w.count("neural_rgb.mlp.linears")
        for i in range(n):
            x = self.w.linear(f"neural_rgb.mlp.linears.{i}", x)
            if i != n - 1:
                x = F.relu(x)
        return x.sigmoid()

    @torch.no_grad()
    def sample_dists(self, center, ray, near, far, rands):
        """model.py:232-248 with the stratified offsets passed in (rands [R,n_coarse] in [0,1))."""
        t = (rands + torch.arange(self.n_coarse)) / self.n_coarse
        dists = t * (far - near) + near                                  # [R,N]
        sdfs = self.sdf_feat(center[:, None] + ray[:, None] * dists[..., None], False)[0][..., 0]
        unif = rr.midpoint_unif(self.n_fine)
        for h in range(self.n_hier):
            fine = rr.neus_upsample(dists, sdfs, unif, 64.0 * 2 ** h)
            s_fine = None
            if h != self.n_hier - 1:
                s_fine = self.sdf_feat(center[:, None] + ray[:, None] * fine[..., None], False)[0][..., 0]
            dists, merged = rr.merge_sorted(dists, sdfs if s_fine is not None else None, fine, s_fine)
            sdfs = merged if merged is not None else sdfs
        return dists

    def render(self, center, ray, rands, progress=1.0):
        """model.py:124-190 (training mode, no background).  center, ray [R,3] (ray unit)."""
        ctc, ctv = (center * center).sum(-1, keepdim=True), (center * ray).sum(-1, keepdim=True)
        disc = (ctv ** 2 - (ctc - 1.0)).sqrt()
        near, far = (-ctv - disc).relu(), -ctv + disc
        outside = near.isnan()
        near = torch.where(outside, torch.ones_like(near), near)
        far = torch.where(outside, torch.full_like(far, 1.2), far)
        dists = self.sample_dists(center, ray, near, far, rands)
        pts = center[:, None] + ray[:, None] * dists[..., None]
        sdf, feat = self.sdf_feat(pts)
        sdf = torch.where(outside[..., None].expand_as(sdf), torch.full_like(sdf, 1000.0), sdf)
        grad, hess = self.gradients(pts, sdf)
        rgbs = self.rgb(pts, F.normalize(grad, dim=-1), ray[:, None].expand_as(pts), feat)
        ratio = min(progress / self.anneal_end, 1.0)
        rgb, weights, _ = rr.neus_render(sdf[..., 0], grad, dists, ray, far[:, 0], self.w.p["s_var"], ratio, rgbs)
        return dict(rgb=rgb, gradients=grad, hessians=hess, outside=outside, dists=dists, weights=weights)

    def loss(self, out, target):
        """trainer.py:40-48 with base.yaml's weights; misc.py:74-89."""
        keep = (~out["outside"]).float()
        eik = (((out["gradients"].norm(dim=-1) - 1.0) ** 2).nan_to_num(0.0, 0.0, 0.0) * keep).mean()
        curv = (out["hessians"].sum(dim=-1).abs().nan_to_num(0.0, 0.0, 0.0) * keep).mean()
        return 3 * F.l1_loss(out["rgb"], target) + 0.1 * eik + 5e-4 * curv

    @torch.no_grad()
    def extract_mesh(self, resolution=256, block_res=128):
        """mesh.py:24-133 on [-1,1]^3 with the C marching cubes; returns (vertices, faces, seconds_sdf, seconds_mc)."""
        intv = 2.0 / resolution
        grid = torch.arange(-1.0, 1.0, intv)
        nb = int(np.ceil(len(grid) / block_res))
        verts, faces, nv, t_sdf, t_mc = [], [], 0, 0.0, 0.0
        for bx in range(nb):
            for by in range(nb):
                for bz in range(nb):
                    ax = [grid[b * block_res:b * block_res + block_res + 1] for b in (bx, by, bz)]
                    xyz = torch.stack(torch.meshgrid(*ax, indexing="ij"), dim=-1)
                    t0 = time.perf_counter()
                    sdf = -self.sdf_feat(xyz.reshape(-1, 3), False)[0].reshape(xyz.shape[:-1])
                    t1 = time.perf_counter()
                    v, f = mcubes_ref.marching_cubes(sdf.numpy(), 0.0)
                    t2 = time.perf_counter()
                    t_sdf, t_mc = t_sdf + t1 - t0, t_mc + t2 - t1
                    if len(v):
                        v = v * intv + xyz[0, 0, 0].numpy()
                        keep = np.linalg.norm(v, axis=-1) < 1.0
                        remap = np.full(len(v), -1)
                        remap[keep] = np.arange(keep.sum())
                        fk = f[keep[f].all(axis=1)]
                        verts.append(v[keep]), faces.append(remap[fk] + nv)
                        nv += int(keep.sum())
        v = np.concatenate(verts) if verts else np.zeros((0, 3))
        f = np.concatenate(faces) if faces else np.zeros((0, 3), dtype=np.int64)
        return v, f, t_sdf, t_mc


def synthetic_rays(n_rays, seed=0):
    """camera centre (0,0,-3), directions normalize(N(0,0.15^2) + (0,0,1))."""
    g = torch.Generator().manual_seed(seed)
    d = torch.randn(n_rays, 3, generator=g) * 0.15 + torch.tensor([0.0, 0.0, 1.0])
    center = torch.tensor([0.0, 0.0, -3.0]).expand(n_rays, 3).contiguous()
    return center, F.normalize(d, dim=-1), torch.rand(n_rays, 64, generator=g), torch.rand(n_rays, 3, generator=g)


def geometric_init_state(hidden=256, fourier_levels=6, rgb_layers=4, seed=0):
    """A reference-format state dict with the geometric (sphere r = 0.5) initialisation of mlp.py:71-84."""
    g = torch.Generator().manual_seed(seed)
    k0 = 3 + 6 * fourier_levels
    sd = {}

    def wn(prefix, w, b):
        sd[prefix + ".weight_v"], sd[prefix + ".weight_g"], sd[prefix + ".bias"] = w, w.norm(dim=1, keepdim=True), b

    w0 = torch.randn(hidden, k0, generator=g) * math.sqrt(2 / hidden)
    w0[:, 3:] = 0.0
    wn("neural_sdf.mlp.linears.0", w0, torch.zeros(hidden))
    wn("neural_sdf.mlp.linears.1", torch.randn(hidden, hidden, generator=g) * math.sqrt(2 / hidden), torch.zeros(hidden))
    sd["neural_sdf.mlp.linear_sdf.weight"] = torch.randn(1, hidden, generator=g) * 1e-4 + math.sqrt(math.pi / hidden)
    sd["neural_sdf.mlp.linear_sdf.bias"] = torch.full((1,), -0.5)
    dims = [3 + 16 + 3 + hidden] + [hidden] * rgb_layers + [3]
    for i, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
        bound = 1 / math.sqrt(a)
        wn(f"neural_rgb.mlp.linears.{i}", (torch.rand(b, a, generator=g) * 2 - 1) * bound,
           torch.zeros(b) if i == len(dims) - 2 else (torch.rand(b, generator=g) * 2 - 1) * bound)
    sd["s_var"] = torch.tensor(3.0)
    return sd


def timed_baseline(n_rays=1024, resolution=256, threads=None):
    """The CPU baseline of BASELINE.json configs[0]: forward + backward of one 1024-ray batch, then extraction."""
    if threads:
        torch.set_num_threads(threads)
    path = CpuPath(geometric_init_state(), requires_grad=True)
    center, ray, rands, target = synthetic_rays(n_rays)
    t0 = time.perf_counter()
    out = path.render(center, ray, rands)
    loss = path.loss(out, target)
    t1 = time.perf_counter()
    loss.backward()
    t2 = time.perf_counter()
    v, f, t_sdf, t_mc = path.extract_mesh(resolution)
    return {"rays": n_rays, "render_fwd_s": t1 - t0, "render_bwd_s": t2 - t1, "rays_per_s": n_rays / (t2 - t0),
            "mesh_resolution": resolution, "mesh_sdf_s": t_sdf, "mesh_mc_s": t_mc, "vertices": len(v), "faces": len(f),
            "lattice_points_per_s": (resolution + 1) ** 3 / max(t_sdf + t_mc, 1e-9), "threads": torch.get_num_threads()}
