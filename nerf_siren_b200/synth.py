"""Synthetic inputs shaped like the reference's datasets (no dataset ships with the reference, no network here).

``blender_rays``: datasets/blender.py:30-42,64-69 + datasets/ray_utils.py:5-50 -- an H x W pinhole camera
(camera_angle_x = 0.6911 rad as in the public lego transforms) on a sphere of radius ~4.03 looking at the origin,
ray directions rotated to world space and NORMALISED, near 2 / far 6, white background.
``llff_rays``: datasets/llff.py:235-250 + ray_utils.py:53-94 -- 504 x 378 NDC rays: origins on the z = -1 plane,
un-normalised directions, near 0 / far 1.
Generated with torch on whatever device is asked for (the generator seeds a CPU stream so that every rank / run
sees the same data).  These are inputs only -- no rendering arithmetic lives here.
"""
from __future__ import annotations

import math

import torch


def _camera(g: torch.Generator, radius: float):
    th = torch.rand(1, generator=g).item() * 2 * math.pi
    ph = (0.15 + 0.5 * torch.rand(1, generator=g).item()) * math.pi / 2
    cam = torch.tensor([math.cos(th) * math.cos(ph), math.sin(th) * math.cos(ph), math.sin(ph)]) * radius
    fwd = -cam / cam.norm()
    right = torch.linalg.cross(fwd, torch.tensor([0.0, 0.0, 1.0]))
    right = right / right.norm()
    up = torch.linalg.cross(right, fwd)
    return cam, torch.stack([right, up, -fwd], dim=1)          # camera-to-world rotation, columns x, y, z(back)


def blender_image_rays(h: int = 800, w: int = 800, seed: int = 0, radius: float = 4.0311, near: float = 2.0,
                       far: float = 6.0) -> torch.Tensor:
    """All H*W rays of one synthetic Blender view, row-major like datasets/blender.py:103-108 -> (H*W, 8)."""
    g = torch.Generator().manual_seed(seed)
    cam, c2w = _camera(g, radius)
    focal = 0.5 * w / math.tan(0.5 * 0.6911112070083618)
    j, i = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    d_cam = torch.stack([(i - w / 2) / focal, -(j - h / 2) / focal, -torch.ones_like(i)], dim=-1).reshape(-1, 3)
    d = d_cam @ c2w.T
    d = d / d.norm(dim=-1, keepdim=True)
    n = d.shape[0]
    return torch.cat([cam.expand(n, 3), d, torch.full((n, 1), near), torch.full((n, 1), far)], dim=1).contiguous()


def blender_rays(n: int, seed: int = 0, hw: int = 800, n_views: int = 8) -> torch.Tensor:
    """n rays drawn at random from n_views synthetic Blender views (a training batch) -> (n, 8)."""
    g = torch.Generator().manual_seed(seed)
    out = []
    per = (n + n_views - 1) // n_views
    for v in range(n_views):
        full = blender_image_rays(hw, hw, seed=seed * 1000 + v)
        idx = torch.randint(0, full.shape[0], (per,), generator=g)
        out.append(full[idx])
    rays = torch.cat(out, 0)[:n]
    return rays[torch.randperm(n, generator=g)].contiguous()


def llff_rays(n: int, seed: int = 0, w: int = 504, h: int = 378, focal: float = 407.0) -> torch.Tensor:
    """n NDC rays of an LLFF-like forward-facing capture -> (n, 8)."""
    g = torch.Generator().manual_seed(seed)
    ij = torch.randint(0, w * h, (n,), generator=g)
    i, j = (ij % w).float(), (ij // w).float()
    d = torch.stack([(i - w / 2) / focal, -(j - h / 2) / focal, -torch.ones(n)], dim=-1)
    o = (0.05 * torch.randn(3, generator=g)).expand(n, 3)
    t = -(1.0 + o[:, 2]) / d[:, 2]                                      # ray_utils.py:78-80, near plane 1.0
    o = o + t[:, None] * d
    ox, oy, oz = o[:, 0] / o[:, 2], o[:, 1] / o[:, 2], o[:, 2]
    o_ndc = torch.stack([-focal / (w / 2) * ox, -focal / (h / 2) * oy, 1 + 2.0 / oz], dim=-1)
    d_ndc = torch.stack([-focal / (w / 2) * (d[:, 0] / d[:, 2] - ox), -focal / (h / 2) * (d[:, 1] / d[:, 2] - oy),
                         1 - o_ndc[:, 2]], dim=-1)
    return torch.cat([o_ndc, d_ndc, torch.zeros(n, 1), torch.ones(n, 1)], dim=1).contiguous()


def target_rgbs(n: int, seed: int = 0) -> torch.Tensor:
    return torch.rand(n, 3, generator=torch.Generator().manual_seed(seed + 7919))


# ---- EG3D-shaped inputs (BASELINE config 5) -------------------------------------------------------------------
def eg3d_planes(res: int = 256, channels: int = 32, seed: int = 0) -> torch.Tensor:
    """(3, C, res, res) tri-planes: smooth low-frequency noise plus white noise, so densities vary along a ray (the
    reference's planes come out of a StyleGAN2 backbone, eg3d_training/triplane.py:61, which is out of scope)."""
    g = torch.Generator().manual_seed(seed)
    lo = torch.randn(3, channels, max(res // 16, 2), max(res // 16, 2), generator=g)
    up = torch.nn.functional.interpolate(lo, size=(res, res), mode="bilinear", align_corners=False)
    return (up * 2.0 + 0.3 * torch.randn(3, channels, res, res, generator=g)).contiguous()


def eg3d_decoder(seed: int = 0) -> torch.Tensor:
    """OSGDecoder weights (eg3d_training/triplane.py:144-167, randn init of networks_stylegan2.py:109) as the flat vector
    the kernel takes: W1 (64 x 32) | b1 (64) | W2 (4 x 64) | b2 (4)."""
    g = torch.Generator().manual_seed(seed)
    return torch.cat([torch.randn(64 * 32, generator=g), torch.randn(64, generator=g) * 0.1,
                      torch.randn(4 * 64, generator=g), torch.randn(4, generator=g) * 0.1]).contiguous()


def eg3d_camera_rays(n: int, seed: int = 0, radius: float = 2.7, fov_deg: float = 18.8, res: int = 128):
    """n rays of a res x res pinhole camera on a sphere of `radius` looking at the origin (eg3d_renderer.py:30-32 shapes)
    -> (origins (n, 3), directions (n, 3)), row-major pixel order when n == res * res."""
    g = torch.Generator().manual_seed(seed)
    th, ph = float(torch.rand(1, generator=g)) * 2 * math.pi, 1.2 + 0.6 * float(torch.rand(1, generator=g))
    cam = torch.tensor([radius * math.sin(ph) * math.cos(th), radius * math.cos(ph), radius * math.sin(ph) * math.sin(th)])
    fwd = -cam / cam.norm()
    right = torch.linalg.cross(fwd, torch.tensor([0.0, 1.0, 0.0]))
    right = right / right.norm()
    upv = torch.linalg.cross(right, fwd)
    pix = torch.arange(n) % (res * res) if n >= res * res else torch.randint(0, res * res, (n,), generator=g)
    i, j = (pix % res).float(), (pix // res).float()
    t = math.tan(math.radians(fov_deg) / 2)
    x, y = ((i + 0.5) / res * 2 - 1) * t, (1 - (j + 0.5) / res * 2) * t
    d = fwd[None] + x[:, None] * right[None] + y[:, None] * upv[None]
    d = d / d.norm(dim=-1, keepdim=True)
    return cam.expand(n, 3).contiguous(), d.contiguous()
