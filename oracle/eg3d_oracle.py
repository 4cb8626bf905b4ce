"""CPU oracle for the EG3D tri-plane renderer  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

torch-CPU fp32 restatement of volumetric_rendering/renderer.py (ImportanceRenderer), ray_marcher.py
(MipRayMarcher2) and eg3d_training/triplane.py (OSGDecoder over networks_stylegan2.FullyConnectedLayer),
written functionally: plain tensors in, plain tensors out, and the two random tensors the reference draws
(`torch.rand_like` in sample_stratified, `torch.rand` in sample_pdf) are explicit arguments.

Parity pin: oracle/make_goldens_eg3d.py imports the real reference, asserts bit-equality with this module on
seeded inputs and writes the REFERENCE's outputs to tests/golden/g10_eg3d_*.npz.  Only tests/, smoke() and
bench.py's cpu_baseline leg may import this file.  Citations are file:line into the reference tree.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F


def generate_planes() -> torch.Tensor:
    """renderer.py:23-39 -- note planes 1 and 2 both project onto (x, z)-type axes (upstream quirk, preserved)."""
    return torch.tensor([[[1, 0, 0], [0, 1, 0], [0, 0, 1]],
                         [[1, 0, 0], [0, 0, 1], [0, 1, 0]],
                         [[0, 0, 1], [1, 0, 0], [0, 1, 0]]], dtype=torch.float32)


def project_onto_planes(planes: torch.Tensor, coordinates: torch.Tensor) -> torch.Tensor:
    """renderer.py:41-55."""
    N, M, _ = coordinates.shape
    n_planes = planes.shape[0]
    coordinates = coordinates.unsqueeze(1).expand(-1, n_planes, -1, -1).reshape(N * n_planes, M, 3)
    inv_planes = torch.linalg.inv(planes).unsqueeze(0).expand(N, -1, -1, -1).reshape(N * n_planes, 3, 3)
    return torch.bmm(coordinates, inv_planes)[..., :2]


def sample_from_planes(plane_axes, plane_features, coordinates, box_warp) -> torch.Tensor:
    """renderer.py:57-68: bilinear grid_sample, zeros padding, align_corners=False -> (N, 3, M, C)."""
    N, n_planes, C, H, W = plane_features.shape
    M = coordinates.shape[1]
    plane_features = plane_features.view(N * n_planes, C, H, W)
    coordinates = (2 / box_warp) * coordinates
    proj = project_onto_planes(plane_axes, coordinates).unsqueeze(1)
    out = F.grid_sample(plane_features, proj.float(), mode="bilinear", padding_mode="zeros", align_corners=False)
    return out.permute(0, 3, 2, 1).reshape(N, n_planes, M, C)


def fully_connected(x, weight, bias, lr_multiplier: float = 1.0):
    """networks_stylegan2.py:95-122, activation='linear' branch: addmm(b, x, (w * gain).t())."""
    w = weight * (lr_multiplier / (weight.shape[1] ** 0.5))
    b = bias * lr_multiplier if lr_multiplier != 1 else bias
    return torch.addmm(b.unsqueeze(0), x, w.t())


def osg_decoder(p: Dict[str, torch.Tensor], sampled_features: torch.Tensor, lr_mul: float = 1.0):
    """triplane.py:144-167: mean over planes -> FC(32,64) -> softplus -> FC(64,1+3); sigmoid clamp of MipNeRF."""
    x = sampled_features.mean(1)
    N, M, C = x.shape
    x = x.view(N * M, C)
    x = fully_connected(x, p["net.0.weight"], p["net.0.bias"], lr_mul)
    x = F.softplus(x)
    x = fully_connected(x, p["net.2.weight"], p["net.2.bias"], lr_mul)
    x = x.view(N, M, -1)
    rgb = torch.sigmoid(x[..., 1:]) * (1 + 2 * 0.001) - 0.001
    return rgb, x[..., 0:1]


def run_model(planes, dec, coords, box_warp, lr_mul: float = 1.0):
    """renderer.py:144-151 without density noise."""
    feats = sample_from_planes(generate_planes(), planes, coords, box_warp)
    return osg_decoder(dec, feats, lr_mul)


def mip_ray_march(colors, densities, depths, white_back: bool = False):
    """ray_marcher.py:25-57 (clamp_mode softplus)."""
    deltas = depths[:, :, 1:] - depths[:, :, :-1]
    colors_mid = (colors[:, :, :-1] + colors[:, :, 1:]) / 2
    densities_mid = (densities[:, :, :-1] + densities[:, :, 1:]) / 2
    depths_mid = (depths[:, :, :-1] + depths[:, :, 1:]) / 2
    densities_mid = F.softplus(densities_mid - 1)
    density_delta = densities_mid * deltas
    alpha = 1 - torch.exp(-density_delta)
    alpha_shifted = torch.cat([torch.ones_like(alpha[:, :, :1]), 1 - alpha + 1e-10], -2)
    weights = alpha * torch.cumprod(alpha_shifted, -2)[:, :, :-1]
    composite_rgb = torch.sum(weights * colors_mid, -2)
    weight_total = weights.sum(2)
    composite_depth = torch.sum(weights * depths_mid, -2) / weight_total
    composite_depth = torch.nan_to_num(composite_depth, float("inf"))
    composite_depth = torch.clamp(composite_depth, torch.min(depths), torch.max(depths))
    if white_back:
        composite_rgb = composite_rgb + 1 - weight_total
    return composite_rgb, composite_depth, weights


def sample_stratified(N: int, M: int, ray_start, ray_end, depth_resolution: int, rand: torch.Tensor,
                      disparity_space_sampling: bool = False) -> torch.Tensor:
    """renderer.py:172-195; `rand` (N, M, S, 1) replaces torch.rand_like; ray_start/end floats or (N, M, 1) tensors."""
    S = depth_resolution
    if disparity_space_sampling:
        d = torch.linspace(0, 1, S).reshape(1, 1, S, 1).repeat(N, M, 1, 1)
        d = d + rand * (1 / (S - 1))
        return 1. / (1. / ray_start * (1. - d) + 1. / ray_end * d)
    if isinstance(ray_start, torch.Tensor):
        steps = torch.arange(S, dtype=torch.float32) / (S - 1)
        for _ in range(ray_start.ndim):
            steps = steps.unsqueeze(-1)
        d = (ray_start[None] + steps * (ray_end - ray_start)[None]).permute(1, 2, 0, 3)      # math_utils.py:101-118
        delta = (ray_end - ray_start) / (S - 1)
        return d + rand * delta[..., None]
    d = torch.linspace(ray_start, ray_end, S).reshape(1, 1, S, 1).repeat(N, M, 1, 1)
    return d + rand * ((ray_end - ray_start) / (S - 1))


def sample_pdf(bins, weights, n_importance: int, u: torch.Tensor, eps: float = 1e-5):
    """renderer.py:217-256 with explicit u (N_rays, n_importance); also returns the int64 indices."""
    n_rays, n_samples_ = weights.shape
    weights = weights + eps
    pdf = weights / torch.sum(weights, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[:, :1]), cdf], -1)
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.clamp_min(inds - 1, 0)
    above = torch.clamp_max(inds, n_samples_)
    inds_sampled = torch.stack([below, above], -1).view(n_rays, 2 * n_importance)
    cdf_g = torch.gather(cdf, 1, inds_sampled).view(n_rays, n_importance, 2)
    bins_g = torch.gather(bins, 1, inds_sampled).view(n_rays, n_importance, 2)
    denom = cdf_g[..., 1] - cdf_g[..., 0]
    denom[denom < eps] = 1
    return bins_g[..., 0] + (u - cdf_g[..., 0]) / denom * (bins_g[..., 1] - bins_g[..., 0]), inds, cdf


def smooth_weights(weights2d: torch.Tensor) -> torch.Tensor:
    """renderer.py:207-210: max_pool1d(k2,s1,p1) -> avg_pool1d(k2,s1) -> + 0.01 on (rays, S-1) weights."""
    w = F.max_pool1d(weights2d.unsqueeze(1).float(), 2, 1, padding=1)
    w = F.avg_pool1d(w, 2, 1).squeeze(1)
    return w + 0.01


def sample_importance(z_vals, weights, n_importance: int, u: torch.Tensor):
    """renderer.py:197-215."""
    B, R, S, _ = z_vals.shape
    z = z_vals.reshape(B * R, S)
    w = smooth_weights(weights.reshape(B * R, -1))
    z_mid = 0.5 * (z[:, :-1] + z[:, 1:])
    s, _, _ = sample_pdf(z_mid, w[:, 1:-1], n_importance, u)
    return s.detach().reshape(B, R, n_importance, 1)


def unify_samples(d1, c1, s1, d2, c2, s2):
    """renderer.py:160-170."""
    all_d = torch.cat([d1, d2], dim=-2)
    all_c = torch.cat([c1, c2], dim=-2)
    all_s = torch.cat([s1, s2], dim=-2)
    _, idx = torch.sort(all_d, dim=-2)
    return (torch.gather(all_d, -2, idx), torch.gather(all_c, -2, idx.expand(-1, -1, -1, all_c.shape[-1])),
            torch.gather(all_s, -2, idx.expand(-1, -1, -1, 1)))


def importance_render(planes, dec, ray_origins, ray_directions, opts: dict, rand_coarse: torch.Tensor,
                      u: torch.Tensor, lr_mul: float = 1.0) -> Tuple[torch.Tensor, ...]:
    """renderer.py:94-142 (numeric ray_start / ray_end).  Returns the reference's 6 outputs."""
    B, R, _ = ray_origins.shape
    S, Si = opts["depth_resolution"], opts["depth_resolution_importance"]
    wb = opts.get("white_back", False)
    depths_coarse = sample_stratified(B, R, opts["ray_start"], opts["ray_end"], S, rand_coarse,
                                      opts.get("disparity_space_sampling", False))
    coords = (ray_origins.unsqueeze(-2) + depths_coarse * ray_directions.unsqueeze(-2)).reshape(B, -1, 3)
    rgb, sigma = run_model(planes, dec, coords, opts["box_warp"], lr_mul)
    colors_coarse = rgb.reshape(B, R, S, rgb.shape[-1])
    dens_coarse = sigma.reshape(B, R, S, 1)
    rgb_c, depth_c, w_c = mip_ray_march(colors_coarse, dens_coarse, depths_coarse, wb)
    depths_fine = sample_importance(depths_coarse, w_c, Si, u)
    coords = (ray_origins.unsqueeze(-2) + depths_fine * ray_directions.unsqueeze(-2)).reshape(B, -1, 3)
    rgb, sigma = run_model(planes, dec, coords, opts["box_warp"], lr_mul)
    colors_fine = rgb.reshape(B, R, Si, rgb.shape[-1])
    dens_fine = sigma.reshape(B, R, Si, 1)
    all_d, all_c, all_s = unify_samples(depths_coarse, colors_coarse, dens_coarse, depths_fine, colors_fine, dens_fine)
    rgb_f, depth_f, w = mip_ray_march(all_c, all_s, all_d, wb)
    return rgb_c, depth_c, w_c.sum(2), rgb_f, depth_f, w.sum(2)


# ---- seeded synthetic inputs (shared by goldens, tests, bench) ----------------------------------------------
def decoder_init(seed: int = 0, n_features: int = 32, hidden: int = 64, out: int = 4, gain: float = 1.0) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    return {"net.0.weight": torch.randn(hidden, n_features, generator=g) * gain,
            "net.0.bias": torch.randn(hidden, generator=g) * 0.1,
            "net.2.weight": torch.randn(out, hidden, generator=g) * gain,
            "net.2.bias": torch.randn(out, generator=g) * 0.1}


def planes_init(seed: int = 0, res: int = 256, channels: int = 32, scale: float = 1.0) -> torch.Tensor:
    """Smooth-ish random tri-planes (1, 3, C, res, res): low-res noise upsampled + white noise, so densities vary."""
    g = torch.Generator().manual_seed(seed)
    lo = torch.randn(3, channels, max(res // 16, 2), max(res // 16, 2), generator=g)
    up = F.interpolate(lo, size=(res, res), mode="bilinear", align_corners=False)
    return ((up * 2.0 + 0.3 * torch.randn(3, channels, res, res, generator=g)) * scale).unsqueeze(0).contiguous()


def camera_rays(n: int, seed: int = 0, radius: float = 2.7, fov_deg: float = 18.8, res: int = 128) -> Tuple[torch.Tensor, torch.Tensor]:
    """n rays of a res x res pinhole camera on a sphere of `radius` looking at the origin (EG3D-shaped: eg3d_renderer.py:30-32)."""
    g = torch.Generator().manual_seed(seed)
    th, ph = float(torch.rand(1, generator=g)) * 6.28318, 1.2 + 0.6 * float(torch.rand(1, generator=g))
    import math
    cam = torch.tensor([radius * math.sin(ph) * math.cos(th), radius * math.cos(ph), radius * math.sin(ph) * math.sin(th)])
    fwd = -cam / cam.norm()
    up = torch.tensor([0., 1., 0.])
    right = torch.linalg.cross(fwd, up); right = right / right.norm()
    upv = torch.linalg.cross(right, fwd)
    pix = torch.randint(0, res * res, (n,), generator=g)
    i, j = (pix % res).float(), (pix // res).float()
    t = math.tan(math.radians(fov_deg) / 2)
    x = ((i + 0.5) / res * 2 - 1) * t
    y = (1 - (j + 0.5) / res * 2) * t
    d = fwd[None] + x[:, None] * right[None] + y[:, None] * upv[None]
    d = d / d.norm(dim=-1, keepdim=True)
    return cam[None].expand(n, 3).contiguous().unsqueeze(0), d.unsqueeze(0).contiguous()
