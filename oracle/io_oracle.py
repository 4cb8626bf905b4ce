"""CPU oracle for the steps either side of render_rays  --  TEST INFRASTRUCTURE.

Restates datasets/ray_utils.py:5-94 (ray generation, NDC), eval.py:163-193 (uint8 quantisation, nan_to_num, PSNR),
datasets/depth_utils.py:43-69 (PFM writer) and extract_color_mesh.py:117-141 (dense query grid).  Pinned by
oracle/make_goldens_io.py against the reference's own functions loaded by file path (kornia, which the reference uses
only for create_meshgrid, is absent and stubbed there with its documented semantics: grid[..., 0] = column, [..., 1] = row).
"""
from __future__ import annotations

import io
import sys

import numpy as np
import torch


def get_ray_directions(H, W, focal):
    """ray_utils.py:5-24."""
    j, i = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    return torch.stack([(i - W / 2) / focal, -(j - H / 2) / focal, -torch.ones_like(i)], -1)


def get_rays(directions, c2w):
    """ray_utils.py:27-50."""
    rays_d = directions @ c2w[:, :3].T
    rays_d = rays_d / torch.norm(rays_d, dim=-1, keepdim=True)
    rays_o = c2w[:, 3].expand(rays_d.shape)
    return rays_o.reshape(-1, 3), rays_d.reshape(-1, 3)


def get_ndc_rays(H, W, focal, near, rays_o, rays_d):
    """ray_utils.py:53-94."""
    t = -(near + rays_o[..., 2]) / rays_d[..., 2]
    rays_o = rays_o + t[..., None] * rays_d
    ox_oz = rays_o[..., 0] / rays_o[..., 2]
    oy_oz = rays_o[..., 1] / rays_o[..., 2]
    o0 = -1. / (W / (2. * focal)) * ox_oz
    o1 = -1. / (H / (2. * focal)) * oy_oz
    o2 = 1. + 2. * near / rays_o[..., 2]
    d0 = -1. / (W / (2. * focal)) * (rays_d[..., 0] / rays_d[..., 2] - ox_oz)
    d1 = -1. / (H / (2. * focal)) * (rays_d[..., 1] / rays_d[..., 2] - oy_oz)
    d2 = 1 - o2
    return torch.stack([o0, o1, o2], -1), torch.stack([d0, d1, d2], -1)


def image_rays(c2w, H, W, focal, near, far, ndc=False):
    """datasets/blender.py:103-108 / datasets/llff.py:294-306: the (H*W, 8) ray rows of one view."""
    o, d = get_rays(get_ray_directions(H, W, focal), c2w)
    if ndc:
        near, far = 0, 1
        o, d = get_ndc_rays(H, W, focal, 1.0, o, d)
    return torch.cat([o, d, near * torch.ones_like(o[:, :1]), far * torch.ones_like(o[:, :1])], 1)


def quantize(img: np.ndarray) -> np.ndarray:
    """eval.py:187."""
    return (img * 255).astype(np.uint8)


def psnr(pred: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """metrics.py:4-14."""
    return -10 * torch.log10(torch.mean((pred - gt) ** 2))


def pfm_bytes(image: np.ndarray, scale=1) -> bytes:
    """datasets/depth_utils.py:43-69 into a bytes object."""
    f = io.BytesIO()
    image = np.flipud(image)
    color = len(image.shape) == 3 and image.shape[2] == 3
    f.write(("PF\n" if color else "Pf\n").encode("utf-8"))
    f.write("{} {}\n".format(image.shape[1], image.shape[0]).encode("utf-8"))
    endian = image.dtype.byteorder
    if endian == "<" or (endian == "=" and sys.byteorder == "little"):
        scale = -scale
    f.write(("%f\n" % scale).encode("utf-8"))
    f.write(np.ascontiguousarray(image).tobytes())
    return f.getvalue()


def grid_points(N, x_range, y_range, z_range) -> torch.Tensor:
    """extract_color_mesh.py:117-125."""
    x = np.linspace(x_range[0], x_range[1], N)
    y = np.linspace(y_range[0], y_range[1], N)
    z = np.linspace(z_range[0], z_range[1], N)
    return torch.FloatTensor(np.stack(np.meshgrid(x, y, z), -1).reshape(-1, 3))


def vol_pack(rgbsigma: np.ndarray, xmin: float, xmax: float, N: int) -> np.ndarray:
    """extract_mesh.ipynb cells 4 + 7 (the `.vol` file Unity's volume renderer reads), the notebook's numpy statements on an
    (N^3, 4) float32 [rgb sigma] array; returns the uint32 array whose bytes are the file.  (The notebook asserts N == 512;
    the arithmetic does not depend on it.)"""
    sigma = np.maximum(rgbsigma[:, -1], 0)                      # cell 4
    a = 1 - np.exp(-(xmax - xmin) / N * sigma)
    a = a.flatten()
    rgb = (rgbsigma[:, :3] * 255).astype(np.uint32)
    i = np.where(a > 0)[0]
    rgb = rgb[i]
    a = a[i]
    s = rgb.dot(np.array([1 << 24, 1 << 16, 1 << 8])) + (a * 255).astype(np.uint32)
    return np.stack([i, s], -1).astype(np.uint32).flatten()
