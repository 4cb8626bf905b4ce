"""The output side of the path: what ``eval.py:70-198``, ``datasets/depth_utils.py:43-69`` and
``extract_color_mesh.py:117-144`` do around ``render_rays``, with the per-pixel work in kernels (``rays_image.cu``, the
grid input mode of kernel (1)) and only file writing on the host.
"""
from __future__ import annotations

import sys
from collections import defaultdict
from typing import Optional

import numpy as np
import torch

from . import _field, _lib, ops
from .models.rendering import render_rays


@torch.no_grad()
def batched_inference(models, embeddings, rays, N_samples, N_importance, use_disp, chunk, white_back, render_func=render_rays,
                      **kwargs):
    """eval.py:70-103: inference on rays in chunks of 32k, test_time=True."""
    chunk = 1024 * 32                                        # hard-coded in the reference (eval.py:79)
    results = defaultdict(list)
    for i in range(0, rays.shape[0], chunk):
        out = render_func(models, embeddings, rays[i:i + chunk], N_samples, use_disp, 0, 0, N_importance, chunk, white_back,
                          test_time=True, **kwargs)
        for k, v in out.items():
            results[k] += [v]
    return {k: torch.cat(v, 0) for k, v in results.items()}


def save_pfm(filename, image, scale=1, already_flipped=False):
    """datasets/depth_utils.py:43-69, byte-identical output.  ``already_flipped``: the payload comes from
    ops.finish_image, which has flipped the rows on the device."""
    if image.dtype.name != "float32":
        raise Exception("Image dtype must be float32.")
    if len(image.shape) == 3 and image.shape[2] == 3:
        color = True
    elif len(image.shape) == 2 or (len(image.shape) == 3 and image.shape[2] == 1):
        color = False
    else:
        raise Exception("Image must have H x W x 3, H x W x 1 or H x W dimensions.")
    if not already_flipped:
        image = np.flipud(image)
    with open(filename, "wb") as f:
        f.write(("PF\n" if color else "Pf\n").encode("utf-8"))
        f.write("{} {}\n".format(image.shape[1], image.shape[0]).encode("utf-8"))
        endian = image.dtype.byteorder
        if endian == "<" or (endian == "=" and sys.byteorder == "little"):
            scale = -scale
        f.write(("%f\n" % scale).encode("utf-8"))
        np.ascontiguousarray(image).tofile(f)


@torch.no_grad()
def render_image(models, embeddings, c2w, h, w, focal, near, far, N_samples, N_importance, use_disp=False, white_back=True,
                 ndc=False, gt: Optional[torch.Tensor] = None, device="cuda"):
    """One eval.py:152-193 iteration with nothing per-pixel on the host: rays generated on the device (row f2), rendered
    (test_time=True), quantised / flipped / scored on the device (row f3).
    Returns {'img': uint8 (H,W,3) tensor, 'depth_pfm': float32 (H,W) rows flipped, 'psnr': float or None, 'results': dict}."""
    rays = ops.generate_rays(c2w, h, w, focal, near, far, ndc, device)
    results = batched_inference(models, embeddings, rays, N_samples, N_importance, use_disp, 32768, white_back)
    typ = "fine" if "rgb_fine" in results else "coarse"
    depth = results.get(f"depth_{typ}")
    img, depth_pfm, se = ops.finish_image(results[f"rgb_{typ}"].contiguous(), h, w, None if depth is None else depth.contiguous(),
                                          None if gt is None else gt.reshape(-1, 3).contiguous().to(device))
    p = None
    if se is not None:
        p = float(-10 * torch.log10(se / (3 * h * w)))
    return {"img": img, "depth_pfm": depth_pfm, "psnr": p, "results": results}


@torch.no_grad()
def sigma_grid(nerf, embedding_xyz, N_grid, x_range, y_range, z_range):
    """extract_color_mesh.py:117-141: occupancy max(sigma, 0) on the dense N^3 grid -> (N, N, N) CUDA tensor, in the layout
    ``mcubes.marching_cubes`` receives there.  Points are generated and encoded inside kernel (1); the MLP runs sigma-only
    (sigma does not depend on the direction, extract_color_mesh.py:127)."""
    desc = nerf._desc(False, embedding_xyz.N_freqs, 4, embedding_xyz.logscale)
    if not ops.mlp_supported(desc, _lib.PREC_BF16, False):
        raise RuntimeError("the dense-grid query runs on the tcgen05 arm: this model shape is outside it")
    return ops.sigma_grid(desc, _field.flat_params(nerf).detach(), int(N_grid), x_range, y_range, z_range)


@torch.no_grad()
def export_vol(filename, nerf, embedding_xyz, N_grid, x_range, y_range, z_range):
    """extract_mesh.ipynb cells 4 + 7: the sparse ``.vol`` volume for Unity -- dense (N^3, 4) [rgb sigma] query with view
    direction 0 (kernel (1), grid input), alpha = 1 - exp(-(xmax - xmin) / N * max(sigma, 0)), uint32 pairs
    (index, r<<24 | g<<16 | b<<8 | alpha) of the voxels with alpha > 0.  Everything per-voxel runs on the device; the host
    writes the bytes.  Returns the (n_valid, 2) uint32 CUDA tensor.  The reference asserts N == 512 for the Unity shader."""
    desc = nerf._desc(False, embedding_xyz.N_freqs, 4, embedding_xyz.logscale)
    if not ops.mlp_supported(desc, _lib.PREC_BF16, False):
        raise RuntimeError("the dense-grid query runs on the tcgen05 arm: this model shape is outside it")
    rgbsigma = ops.sigma_grid(desc, _field.flat_params(nerf).detach(), int(N_grid), x_range, y_range, z_range, sigma_only=False)
    pairs = ops.pack_vol(rgbsigma, -(x_range[1] - x_range[0]) / N_grid)
    if filename is not None:
        with open(filename, "wb") as f:
            f.write(pairs.cpu().numpy().tobytes())
    return pairs
