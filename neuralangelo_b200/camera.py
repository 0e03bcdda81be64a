"""Camera helpers on the path: `projects/nerf/utils/camera.py:243-320` (get_center_and_ray,
get_3D_points_from_dist) plus `rays_for_indices`, which evaluates only the sampled pixels instead of
all H*W rays followed by a slice (model.py:117-119 builds 1.92 M rays per step to keep 512)."""
import torch


def _pose_invert(pose):
    """camera.py Pose.invert: [R|t]^-1 = [R^T | -R^T t] for [...,3,4]."""
    R, t = pose[..., :3], pose[..., 3:]
    R_inv = R.transpose(-1, -2)
    return torch.cat([R_inv, -R_inv @ t], dim=-1)


def _inv3x3(m):
    """Closed-form inverse of [...,3,3] matrices (adjugate / determinant).  `torch.inverse` goes through a
    LAPACK-style factorisation with a host-side status check, which stalls the stream every step and cannot be
    captured in a CUDA graph."""
    a, b, c = m[..., 0, 0], m[..., 0, 1], m[..., 0, 2]
    d, e, f = m[..., 1, 0], m[..., 1, 1], m[..., 1, 2]
    g, h, i = m[..., 2, 0], m[..., 2, 1], m[..., 2, 2]
    A, B, C = e * i - f * h, c * h - b * i, b * f - c * e
    D, E, F = f * g - d * i, a * i - c * g, c * d - a * f
    G, H, I = d * h - e * g, b * g - a * h, a * e - b * d
    det = a * A + b * D + c * G
    adj = torch.stack([torch.stack([A, B, C], dim=-1), torch.stack([D, E, F], dim=-1), torch.stack([G, H, I], dim=-1)],
                      dim=-2)
    return adj / det[..., None, None]


def _to_hom(X):
    return torch.cat([X, torch.ones_like(X[..., :1])], dim=-1)


def _rays_from_pixels(xy, pose, intr):
    """xy [B,R,2] pixel centres -> (center [B,R,3], ray [B,R,3] at depth 1), camera.py:300-311."""
    grid_3D = _to_hom(xy) @ _inv3x3(intr).transpose(-1, -2)             # img2cam
    pose_inv = _pose_invert(pose)
    center_3D = _to_hom(torch.zeros_like(grid_3D)) @ pose_inv.transpose(-1, -2)   # cam2world of the origin
    grid_3D = _to_hom(grid_3D) @ pose_inv.transpose(-1, -2)
    return center_3D, grid_3D - center_3D


def get_center_and_ray(pose, intr, image_size):
    """camera.py:283-311 for batched pose [B,3,4] / intr [B,3,3]: all H*W rays, row-major pixels."""
    H, W = image_size
    with torch.no_grad():
        y = torch.arange(H, dtype=torch.float32, device=pose.device).add_(0.5)
        x = torch.arange(W, dtype=torch.float32, device=pose.device).add_(0.5)
        Y, X = torch.meshgrid(y, x, indexing="ij")
        xy = torch.stack([X, Y], dim=-1).view(-1, 2)
    xy = xy.repeat(len(pose), 1, 1)
    return _rays_from_pixels(xy, pose, intr)


def rays_for_indices(pose, intr, image_size, ray_idx):
    """Same values as slice_by_ray_idx(get_center_and_ray(...), ray_idx) (model.py:117-119), computed for the
    sampled pixels only.  ray_idx [B,R] indexes the row-major H*W grid."""
    H, W = image_size
    with torch.no_grad():
        xy = torch.stack([(ray_idx % W).float() + 0.5, torch.div(ray_idx, W, rounding_mode="floor").float() + 0.5], dim=-1)
    return _rays_from_pixels(xy, pose, intr)


def ray_setup(pose, intr, image_size, ray_idx, with_bounds=True):
    """rays_for_indices + normalize + the unit-sphere bounds of Model.get_dist_bounds in ONE launch (na_ray_setup) for
    CUDA inputs that need no gradient (the reference never differentiates the cameras on this path).  Returns
    (center, ray, ray_unit [B,R,3], (near, far [B,R,1], outside [B,R,1] bool) | None)."""
    from . import _lib
    if not pose.is_cuda:
        raise _lib.NaError("ray_setup is the CUDA path; use rays_for_indices on the CPU")
    if torch.is_grad_enabled() and (pose.requires_grad or intr.requires_grad):
        raise NotImplementedError("ray_setup has no backward: differentiate rays_for_indices instead")
    H, W = image_size
    B, R = ray_idx.shape
    pose_c, intr_c = _lib.f32c(pose.detach()), _lib.f32c(intr.detach())
    idx = ray_idx.to(torch.int64).contiguous()
    out = torch.empty(3, B, R, 3, dtype=torch.float32, device=pose.device)
    bounds = torch.empty(2, B, R, 1, dtype=torch.float32, device=pose.device) if with_bounds else None
    outside = torch.empty(B, R, 1, dtype=torch.uint8, device=pose.device) if with_bounds else None
    rc = _lib.lib().na_ray_setup(_lib.ptr(pose_c), _lib.ptr(intr_c), _lib.ptr(idx), B, R, int(W), _lib.ptr(out[0]),
                                 _lib.ptr(out[1]), _lib.ptr(out[2]), _lib.ptr(bounds[0]) if with_bounds else None,
                                 _lib.ptr(bounds[1]) if with_bounds else None, _lib.ptr(outside), _lib.current_stream())
    _lib.check(rc, "na_ray_setup")
    return out[0], out[1], out[2], ((bounds[0], bounds[1], outside.view(torch.bool)) if with_bounds else None)


def get_3D_points_from_dist(center, ray_unit, dist, multi=True):
    """camera.py:314-320: x = c + d * v."""
    if multi:
        center, ray_unit = center[..., None, :], ray_unit[..., None, :]
    return center + ray_unit * dist
