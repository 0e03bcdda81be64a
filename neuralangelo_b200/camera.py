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


def get_3D_points_from_dist(center, ray_unit, dist, multi=True):
    """camera.py:314-320: x = c + d * v."""
    if multi:
        center, ray_unit = center[..., None, :], ray_unit[..., None, :]
    return center + ray_unit * dist
