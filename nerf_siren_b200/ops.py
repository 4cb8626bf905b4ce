"""Thin tensor-level wrappers over the C ABI: argument validation, output allocation, current stream.

Every function here launches hand-written sm_100a kernels from libnerf_b200.so on
``torch.cuda.current_stream()``; nothing falls back to torch ops or to the CPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import lib, check, MlpDesc


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _req(t: Optional[torch.Tensor], name: str, dtype=torch.float32, allow_none=False) -> int:
    if t is None:
        if allow_none:
            return 0
        raise ValueError(f"{name} is required")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (nerf_siren_b200 has no CPU path); got {t.device}")
    if t.dtype != dtype:
        raise ValueError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t.data_ptr()


def launch_count() -> int:
    return int(lib.nerfb_launch_count())


def sample_coarse(rays, n_samples, use_disp=False, perturb=0.0, perturb_rand=None):
    n = rays.shape[0]
    if rays.dim() != 2 or rays.shape[1] != 8:
        raise ValueError(f"rays must be (N,8), got {tuple(rays.shape)}")
    z = torch.empty(n, n_samples, device=rays.device, dtype=torch.float32)
    if n == 0:
        return z
    pr = _req(perturb_rand, "perturb_rand", allow_none=not perturb > 0)
    if perturb > 0 and tuple(perturb_rand.shape) != (n, n_samples):
        raise ValueError("perturb_rand must be (N, n_samples)")
    check(lib.nerfb_sample_coarse(_req(rays, "rays"), pr, float(perturb), int(bool(use_disp)), n, n_samples,
                                  z.data_ptr(), _stream()), "nerfb_sample_coarse")
    return z


def embed(x, n_freqs, logscale=True):
    if x.dim() != 2:
        raise ValueError("embed expects (B, C)")
    b, c = x.shape
    out = torch.empty(b, c * (1 + 2 * n_freqs), device=x.device, dtype=torch.float32)
    if b == 0:
        return out
    check(lib.nerfb_embed(_req(x, "x"), b, c, n_freqs, int(bool(logscale)), out.data_ptr(), _stream()), "nerfb_embed")
    return out


def composite_fwd(raw, z, rays, noise, noise_std, white_back):
    """raw (N,S,C) C in {1} u [4,32]; returns (weights, rgb, depth, opacity, extra) -- rgb.. None for C == 1."""
    n, s, c = raw.shape
    dev = raw.device
    weights = torch.empty(n, s, device=dev, dtype=torch.float32)
    opacity = torch.empty(n, device=dev, dtype=torch.float32)
    rgb = depth = extra = None
    if c != 1:
        rgb = torch.empty(n, 3, device=dev, dtype=torch.float32)
        depth = torch.empty(n, device=dev, dtype=torch.float32)
        if c > 4:
            extra = torch.empty(n, c - 4, device=dev, dtype=torch.float32)
    check(lib.nerfb_composite_fwd(_req(raw, "raw"), c, _req(z, "z"), _req(rays, "rays"),
                                  _req(noise, "noise", allow_none=True), float(noise_std), int(bool(white_back)), n, s,
                                  weights.data_ptr(), rgb.data_ptr() if rgb is not None else 0,
                                  depth.data_ptr() if depth is not None else 0, opacity.data_ptr(),
                                  extra.data_ptr() if extra is not None else 0, _stream()), "nerfb_composite_fwd")
    return weights, rgb, depth, opacity, extra


def composite_bwd(raw, z, rays, noise, noise_std, white_back, d_rgb=None, d_depth=None, d_opacity=None,
                  d_weights=None, d_extra=None):
    n, s, c = raw.shape
    d_raw = torch.empty_like(raw)
    check(lib.nerfb_composite_bwd(_req(raw, "raw"), c, _req(z, "z"), _req(rays, "rays"),
                                  _req(noise, "noise", allow_none=True), float(noise_std), int(bool(white_back)), n, s,
                                  _req(d_rgb, "d_rgb", allow_none=True), _req(d_depth, "d_depth", allow_none=True),
                                  _req(d_opacity, "d_opacity", allow_none=True),
                                  _req(d_weights, "d_weights", allow_none=True),
                                  _req(d_extra, "d_extra", allow_none=True), d_raw.data_ptr(), _stream()),
          "nerfb_composite_bwd")
    return d_raw


def searchsorted(a, v, side="left", out=None):
    """torchsearchsorted.searchsorted(a, v, out, side) (src/torchsearchsorted/searchsorted.py:20-53)."""
    if a.dim() != 2 or v.dim() != 2:
        raise ValueError("input `a` and `v` must be 2-D")
    if not (a.shape[0] == v.shape[0] or a.shape[0] == 1 or v.shape[0] == 1):
        raise ValueError("`a` and `v` must have the same number of rows or one of them must have only one")
    if a.device != v.device:
        raise ValueError("`a` and `v` must be on the same device")
    shape = (max(a.shape[0], v.shape[0]), v.shape[1])
    if out is None:
        out = torch.empty(shape, device=v.device, dtype=torch.long)
    elif out.dtype != torch.long or tuple(out.shape) != shape or out.device != a.device:
        raise ValueError("out must be a torch.long tensor of the result shape on the same device")
    if a.dtype != v.dtype or a.dtype not in _SS_DTYPES:
        raise ValueError(f"`a` and `v` must share one of {sorted(str(d) for d in _SS_DTYPES)}; got {a.dtype}, {v.dtype}")
    check(lib.nerfb_searchsorted_typed(_req(a, "a", dtype=a.dtype), a.shape[0], a.shape[1], _req(v, "v", dtype=a.dtype),
                                       v.shape[0], v.shape[1], _SS_DTYPES[a.dtype], 1 if side == "left" else 0,
                                       _req(out, "out", dtype=torch.long), _stream()), "nerfb_searchsorted_typed")
    return out


# the scalar types of the reference's AT_DISPATCH_ALL_TYPES (searchsorted_cuda_kernel.cu:134) -> NERFB_DT_*
_SS_DTYPES = {torch.float32: 0, torch.float64: 1, torch.int64: 2, torch.int32: 3, torch.int16: 4, torch.int8: 5, torch.uint8: 6}


def sample_pdf(bins, weights, n_importance, det=False, eps=1e-5, u=None, return_aux=False):
    n, nb = bins.shape
    if tuple(weights.shape) != (n, nb - 1):
        raise ValueError("weights must be (N, bins-1)")
    if not det and u is None:
        u = torch.rand(n, n_importance, device=bins.device)          # models/rendering.py:47
    samples = torch.empty(n, n_importance, device=bins.device, dtype=torch.float32)
    cdf = inds = None
    if return_aux:
        cdf = torch.empty(n, nb, device=bins.device, dtype=torch.float32)
        inds = torch.empty(n, n_importance, device=bins.device, dtype=torch.long)
    check(lib.nerfb_sample_pdf(_req(bins, "bins"), _req(weights, "weights"), _req(u, "u", allow_none=det),
                               int(bool(det)), float(eps), n, nb, n_importance, samples.data_ptr(),
                               cdf.data_ptr() if cdf is not None else 0, inds.data_ptr() if inds is not None else 0,
                               _stream()), "nerfb_sample_pdf")
    if return_aux:
        return samples, cdf, inds
    return samples


def resample(z_coarse, weights_coarse, n_importance, det, u=None):
    n, sc = z_coarse.shape
    z_fine = torch.empty(n, sc + n_importance, device=z_coarse.device, dtype=torch.float32)
    check(lib.nerfb_resample(_req(z_coarse, "z_coarse"), _req(weights_coarse, "weights_coarse"),
                             _req(u, "u", allow_none=det), int(bool(det)), n, sc, n_importance, z_fine.data_ptr(),
                             _stream()), "nerfb_resample")
    return z_fine


def adam_step(params, grads, exp_avg, exp_avg_sq, step, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, grad_scale=1.0):
    """In-place Adam update of a flat fp32 parameter buffer (one launch); see nerfb_adam_step."""
    n = params.numel()
    for t, name in ((params, "params"), (grads, "grads"), (exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq")):
        if t.numel() != n:
            raise ValueError(f"{name} must have {n} elements")
    check(lib.nerfb_adam_step(_req(params, "params"), _req(grads, "grads"), _req(exp_avg, "exp_avg"),
                              _req(exp_avg_sq, "exp_avg_sq"), n, float(lr), float(betas[0]), float(betas[1]),
                              float(eps), int(step), float(grad_scale), _stream()), "nerfb_adam_step")


# ---------------------------------------------------------------------------------------------------------
# MLP
# ---------------------------------------------------------------------------------------------------------
def mlp_supported(desc: MlpDesc, precision: int, need_grad: bool = False) -> bool:
    """Does this arm implement the model (forward, and backward when need_grad)?"""
    return int(lib.nerfb_mlp_workspace_bytes(C.byref(desc), precision, 128, 0, int(bool(need_grad)))) >= 0


def mlp_workspace(desc: MlpDesc, precision: int, n_points: int, sigma_only: bool, save: bool, device) -> torch.Tensor:
    nbytes = int(lib.nerfb_mlp_workspace_bytes(C.byref(desc), precision, n_points, int(sigma_only), int(save)))
    if nbytes < 0:
        check(nbytes, "nerfb_mlp_workspace_bytes")
    # uint8 storage from the caching allocator is >= 512-byte aligned; round the base up to 1024 ourselves
    buf = torch.empty(nbytes + 1024, device=device, dtype=torch.uint8)
    return buf


def _aligned(buf: torch.Tensor):
    base = buf.data_ptr()
    off = (-base) % 1024
    return base + off, buf.numel() - off


def mlp_fwd(desc, precision, in_mode, params, film, in0, in1, n_rows, n_samples, sigma_only, save, workspace):
    p = n_rows * n_samples if in_mode == _lib.IN_RAYS else n_rows
    c = 1 if sigma_only else int(lib.nerfb_out_channels(C.byref(desc)))
    out = torch.empty(p, c, device=params.device, dtype=torch.float32)
    ws_ptr, ws_bytes = _aligned(workspace)
    check(lib.nerfb_mlp_fwd(C.byref(desc), precision, in_mode, _req(params, "params"), _req(film, "film", allow_none=True),
                            _req(in0, "in0"), _req(in1, "in1", allow_none=in_mode == _lib.IN_EMBEDDED), n_rows, n_samples,
                            int(sigma_only), int(save), out.data_ptr(), ws_ptr, ws_bytes, _stream()), "nerfb_mlp_fwd")
    return out


def mlp_bwd(desc, precision, in_mode, params, film, in0, in1, n_rows, n_samples, out, d_out, workspace, want_d_in=False):
    grad = torch.zeros_like(params)
    d_in = None
    if want_d_in:
        d_in = torch.zeros_like(in0)
    ws_ptr, ws_bytes = _aligned(workspace)
    check(lib.nerfb_mlp_bwd(C.byref(desc), precision, in_mode, _req(params, "params"), _req(film, "film", allow_none=True),
                            _req(in0, "in0"), _req(in1, "in1", allow_none=in_mode == _lib.IN_EMBEDDED), n_rows, n_samples,
                            _req(out, "out"), _req(d_out, "d_out"), grad.data_ptr(),
                            d_in.data_ptr() if d_in is not None else 0, ws_ptr, ws_bytes, _stream()), "nerfb_mlp_bwd")
    return grad, d_in


def render_fused(desc_c, desc_f, precision, params_c, film_c, params_f, film_f, rays, n_samples, n_importance, use_disp,
                 perturb, noise_std, white_back, test_time, perturb_rand=None, noise_coarse=None, u=None, noise_fine=None):
    """The whole inference render_rays in ONE library call (nerfb_render_fused_fwd): returns the tuple
    (rgb_c, depth_c, opacity_c, cls_c, rgb_f, depth_f, opacity_f, cls_f), entries the configuration does not produce = None."""
    n = rays.shape[0]
    dev = rays.device
    cfg = _lib.RenderCfg(precision=int(precision), n_samples=int(n_samples), n_importance=int(n_importance),
                         use_disp=int(bool(use_disp)), white_back=int(bool(white_back)), test_time=int(bool(test_time)),
                         perturb=float(perturb), noise_std=float(noise_std))
    fine = n_importance > 0
    dfp = C.byref(desc_f) if fine else None
    nbytes = int(lib.nerfb_render_fused_workspace_bytes(C.byref(desc_c), dfp, C.byref(cfg), n))
    if nbytes < 0:
        check(nbytes, "nerfb_render_fused_workspace_bytes")
    ws = torch.empty(nbytes + 1024, device=dev, dtype=torch.uint8)
    ws_ptr, ws_bytes = _aligned(ws)
    n_cls_c = int(lib.nerfb_out_channels(C.byref(desc_c))) - 4
    n_cls_f = int(lib.nerfb_out_channels(C.byref(desc_f))) - 4 if fine else 0

    def new(*shape):
        return torch.empty(*shape, device=dev, dtype=torch.float32)
    rgb_c = depth_c = cls_c = rgb_f = depth_f = op_f = cls_f = None
    op_c = new(n)
    if not test_time:
        rgb_c, depth_c = new(n, 3), new(n)
        cls_c = new(n, n_cls_c) if n_cls_c > 0 else None
    if fine:
        rgb_f, depth_f, op_f = new(n, 3), new(n), new(n)
        cls_f = new(n, n_cls_f) if n_cls_f > 0 else None

    def ptr(t):
        return 0 if t is None else t.data_ptr()
    io = _lib.RenderIO(rays=_req(rays, "rays"), perturb_rand=_req(perturb_rand, "perturb_rand", allow_none=not perturb > 0),
                       noise_coarse=_req(noise_coarse, "noise_coarse", allow_none=noise_std == 0),
                       u=_req(u, "u", allow_none=not (perturb > 0 and fine)),
                       noise_fine=_req(noise_fine, "noise_fine", allow_none=(noise_std == 0 or not fine)),
                       rgb_coarse=ptr(rgb_c), depth_coarse=ptr(depth_c), opacity_coarse=ptr(op_c), cls_coarse=ptr(cls_c),
                       rgb_fine=ptr(rgb_f), depth_fine=ptr(depth_f), opacity_fine=ptr(op_f), cls_fine=ptr(cls_f))
    check(lib.nerfb_render_fused_fwd(C.byref(desc_c), dfp, C.byref(cfg), _req(params_c, "params_coarse"),
                                     _req(film_c, "film_coarse", allow_none=True),
                                     _req(params_f, "params_fine", allow_none=not fine),
                                     _req(film_f, "film_fine", allow_none=True), C.byref(io), n, ws_ptr, ws_bytes, _stream()),
          "nerfb_render_fused_fwd")
    return rgb_c, depth_c, op_c, cls_c, rgb_f, depth_f, op_f, cls_f


# ---- kernel (6): EG3D tri-plane renderer ------------------------------------------------------------------
def triplane_relayout(planes):
    """(3, 32, H, W) reference layout -> channels-last (3, H, W, 32)."""
    if planes.dim() != 4 or planes.shape[0] != 3:
        raise ValueError(f"planes must be (3, C, H, W), got {tuple(planes.shape)}")
    _, c, h, w = planes.shape
    out = torch.empty(3, h, w, c, device=planes.device, dtype=torch.float32)
    check(lib.nerfb_triplane_relayout(_req(planes, "planes"), c, h, w, out.data_ptr(), _stream()), "nerfb_triplane_relayout")
    return out


def _check_planes_cl(planes_cl):
    if planes_cl.dim() != 4 or planes_cl.shape[0] != 3 or planes_cl.shape[3] != 32:
        raise ValueError(f"planes_cl must be (3, H, W, 32), got {tuple(planes_cl.shape)}")
    return planes_cl.shape[1], planes_cl.shape[2]


def triplane_field(planes_cl, decoder_flat, lr_mul, coords, box_warp):
    h, w = _check_planes_cl(planes_cl)
    if decoder_flat.numel() != 64 * 32 + 64 + 4 * 64 + 4:
        raise ValueError("decoder_flat must hold W1(64x32) b1(64) W2(4x64) b2(4)")
    if coords.dim() != 2 or coords.shape[1] != 3:
        raise ValueError("coords must be (P, 3)")
    p = coords.shape[0]
    rgb = torch.empty(p, 3, device=coords.device, dtype=torch.float32)
    sigma = torch.empty(p, device=coords.device, dtype=torch.float32)
    if p == 0:
        return rgb, sigma
    check(lib.nerfb_triplane_field(_req(planes_cl, "planes_cl"), h, w, _req(decoder_flat, "decoder"), float(lr_mul),
                                   _req(coords, "coords"), p, float(box_warp), rgb.data_ptr(), sigma.data_ptr(), _stream()),
          "nerfb_triplane_field")
    return rgb, sigma


def mip_march(colors, densities, depths, white_back=False):
    if depths.dim() != 2 or colors.shape != (*depths.shape, 3) or densities.shape != depths.shape:
        raise ValueError("mip_march expects colors (R,S,3), densities (R,S), depths (R,S)")
    r, s = depths.shape
    rgb = torch.empty(r, 3, device=depths.device, dtype=torch.float32)
    depth = torch.empty(r, device=depths.device, dtype=torch.float32)
    weights = torch.empty(r, s - 1, device=depths.device, dtype=torch.float32)
    ws = torch.empty(4, device=depths.device, dtype=torch.int32)
    if r == 0:
        return rgb, depth, weights
    check(lib.nerfb_mip_march(_req(colors, "colors"), _req(densities, "densities"), _req(depths, "depths"), r, s,
                              int(bool(white_back)), rgb.data_ptr(), depth.data_ptr(), weights.data_ptr(), ws.data_ptr(),
                              _stream()), "nerfb_mip_march")
    return rgb, depth, weights


def triplane_field_bwd(planes_cl, decoder_flat, lr_mul, coords, box_warp, d_rgb, d_sigma):
    """Backward of triplane_field: returns (d_planes_cl (3,H,W,32), d_decoder_flat (2372,))."""
    h, w = _check_planes_cl(planes_cl)
    p = coords.shape[0]
    d_planes = torch.zeros_like(planes_cl)
    d_dec = torch.zeros_like(decoder_flat)
    if p == 0:
        return d_planes, d_dec
    check(lib.nerfb_triplane_field_bwd(_req(planes_cl, "planes_cl"), h, w, _req(decoder_flat, "decoder"), float(lr_mul),
                                       _req(coords, "coords"), p, float(box_warp), _req(d_rgb, "d_rgb"), _req(d_sigma, "d_sigma"),
                                       d_planes.data_ptr(), d_dec.data_ptr(), _stream()), "nerfb_triplane_field_bwd")
    return d_planes, d_dec


def mip_march_bwd(colors, densities, depths, white_back, d_rgb=None, d_depth=None, d_weights=None):
    """Backward of mip_march: returns (d_colors (R,S,3), d_densities (R,S))."""
    r, s = depths.shape
    d_colors = torch.zeros(r, s, 3, device=depths.device, dtype=torch.float32)
    d_dens = torch.zeros(r, s, device=depths.device, dtype=torch.float32)
    if r == 0:
        return d_colors, d_dens
    check(lib.nerfb_mip_march_bwd(_req(colors, "colors"), _req(densities, "densities"), _req(depths, "depths"), r, s,
                                  int(bool(white_back)), _req(d_rgb, "d_rgb", allow_none=True), _req(d_depth, "d_depth", allow_none=True),
                                  _req(d_weights, "d_weights", allow_none=True), d_colors.data_ptr(), d_dens.data_ptr(), _stream()),
          "nerfb_mip_march_bwd")
    return d_colors, d_dens


def triplane_samples_bwd(samples, rays_o, rays_d, n_samples, n_importance, white_back, d_rgb_c=None, d_depth_c=None, d_opacity_c=None,
                         d_rgb_f=None, d_depth_f=None, d_opacity_f=None):
    """Backward of the renderer down to its samples: -> (coords (R*St, 3), d_rgb (R*St, 3), d_sigma (R*St,)), the inputs of
    triplane_field_bwd.  ``samples`` (R, 5, St) as returned by triplane_render(..., return_samples=True)."""
    r, st = samples.shape[0], n_samples + n_importance
    if tuple(samples.shape) != (r, 5, st):
        raise ValueError(f"samples must be (R, 5, {st})")
    dev = samples.device
    coords, d_rgb, d_sigma = torch.empty(r * st, 3, device=dev), torch.empty(r * st, 3, device=dev), torch.empty(r * st, device=dev)
    if r == 0:
        return coords, d_rgb, d_sigma
    check(lib.nerfb_triplane_samples_bwd(
        _req(samples, "samples"), _req(rays_o, "rays_o"), _req(rays_d, "rays_d"), r, int(n_samples), int(n_importance), int(bool(white_back)),
        _req(d_rgb_c, "d_rgb_c", allow_none=True), _req(d_depth_c, "d_depth_c", allow_none=True), _req(d_opacity_c, "d_opacity_c", allow_none=True),
        _req(d_rgb_f, "d_rgb_f", allow_none=True), _req(d_depth_f, "d_depth_f", allow_none=True), _req(d_opacity_f, "d_opacity_f", allow_none=True),
        coords.data_ptr(), d_rgb.data_ptr(), d_sigma.data_ptr(), _stream()), "nerfb_triplane_samples_bwd")
    return coords, d_rgb, d_sigma


def triplane_render(planes_cl, decoder_flat, lr_mul, rays_o, rays_d, n_samples, n_importance, ray_start, ray_end,
                    ray_start_arr=None, ray_end_arr=None, disparity=False, box_warp=1.0, white_back=False,
                    rand_coarse=None, u=None, seed=0, return_depths_fine=False, return_samples=False):
    """-> (rgb_c, depth_c, opacity_c, rgb_f, depth_f, opacity_f[, depths_fine (R, Si)] [, samples (R, 5, S + Si)]).
    ``return_samples`` selects nerfb_triplane_render_train (the forward of a training step)."""
    if return_depths_fine and return_samples:
        raise ValueError("samples already hold the importance depths: samples[:, 0, n_samples:]")
    h, w = _check_planes_cl(planes_cl)
    if rays_o.dim() != 2 or rays_o.shape[1] != 3 or rays_d.shape != rays_o.shape:
        raise ValueError("rays_o / rays_d must be (R, 3)")
    r = rays_o.shape[0]
    if rand_coarse is not None and tuple(rand_coarse.shape) != (r, n_samples):
        raise ValueError("rand_coarse must be (R, n_samples)")
    if u is not None and tuple(u.shape) != (r, n_importance):
        raise ValueError("u must be (R, n_importance)")
    dev = rays_o.device
    out = [torch.empty(r, 3, device=dev), torch.empty(r, device=dev), torch.empty(r, device=dev),
           torch.empty(r, 3, device=dev), torch.empty(r, device=dev), torch.empty(r, device=dev)]
    fine = torch.empty(r, n_importance, device=dev) if return_depths_fine else None
    ws = torch.empty(4, device=dev, dtype=torch.int32)
    if return_samples:
        samples = torch.empty(r, 5, n_samples + n_importance, device=dev)
        if r:
            check(lib.nerfb_triplane_render_train(
                _req(planes_cl, "planes_cl"), h, w, _req(decoder_flat, "decoder"), float(lr_mul), _req(rays_o, "rays_o"),
                _req(rays_d, "rays_d"), r, int(n_samples), int(n_importance), float(ray_start), float(ray_end),
                _req(ray_start_arr, "ray_start_arr", allow_none=True), _req(ray_end_arr, "ray_end_arr", allow_none=True),
                int(bool(disparity)), float(box_warp), int(bool(white_back)), _req(rand_coarse, "rand_coarse", allow_none=True),
                _req(u, "u", allow_none=True), int(seed) & 0xFFFFFFFFFFFFFFFF, *[o.data_ptr() for o in out],
                samples.data_ptr(), ws.data_ptr(), _stream()), "nerfb_triplane_render_train")
        return (*out, samples)
    if r == 0:
        return (*out, fine) if return_depths_fine else tuple(out)
    check(lib.nerfb_triplane_render(
        _req(planes_cl, "planes_cl"), h, w, _req(decoder_flat, "decoder"), float(lr_mul), _req(rays_o, "rays_o"),
        _req(rays_d, "rays_d"), r, int(n_samples), int(n_importance), float(ray_start), float(ray_end),
        _req(ray_start_arr, "ray_start_arr", allow_none=True), _req(ray_end_arr, "ray_end_arr", allow_none=True),
        int(bool(disparity)), float(box_warp), int(bool(white_back)), _req(rand_coarse, "rand_coarse", allow_none=True),
        _req(u, "u", allow_none=True), int(seed) & 0xFFFFFFFFFFFFFFFF, *[o.data_ptr() for o in out],
        0 if fine is None else fine.data_ptr(), ws.data_ptr(), _stream()), "nerfb_triplane_render")
    return (*out, fine) if return_depths_fine else tuple(out)


# ---- the steps either side of the path --------------------------------------------------------
def generate_rays(c2w, h, w, focal, near=2.0, far=6.0, ndc=False, device="cuda"):
    """(H*W, 8) rays of one view on the device: ray_utils.get_ray_directions + get_rays (+ get_ndc_rays when ndc)."""
    c2w = torch.as_tensor(c2w, dtype=torch.float32, device="cpu").reshape(3, 4).contiguous()
    rays = torch.empty(h * w, 8, device=device, dtype=torch.float32)
    check(lib.nerfb_generate_rays(c2w.data_ptr(), int(h), int(w), float(focal), float(near), float(far), int(bool(ndc)),
                                  rays.data_ptr(), _stream()), "nerfb_generate_rays")
    return rays


def finish_image(rgb, h, w, depth=None, gt=None):
    """eval.py:163-193 on the device -> (uint8 image (H,W,3), PFM-ordered depth (H,W) or None, squared-error sum or None)."""
    if rgb.shape != (h * w, 3):
        raise ValueError(f"rgb must be ({h * w}, 3), got {tuple(rgb.shape)}")
    u8 = torch.empty(h, w, 3, device=rgb.device, dtype=torch.uint8)
    dp = torch.empty(h, w, device=rgb.device, dtype=torch.float32) if depth is not None else None
    se = torch.zeros(1, device=rgb.device, dtype=torch.float64) if gt is not None else None
    check(lib.nerfb_finish_image(_req(rgb, "rgb"), _req(depth, "depth", allow_none=True), _req(gt, "gt", allow_none=True), int(h),
                                 int(w), u8.data_ptr(), 0 if dp is None else dp.data_ptr(), 0 if se is None else se.data_ptr(),
                                 _stream()), "nerfb_finish_image")
    return u8, dp, se


def sigma_grid(desc, params, n_grid, x_range, y_range, z_range, sigma_only=True):
    """Dense query grid of extract_color_mesh.py:117-144 / extract_mesh.ipynb cell 4, points generated in-kernel:
    sigma_only -> (N, N, N) max(sigma, 0); else (N^3, 4) [rgb sigma] with view direction 0."""
    dev = params.device
    rg = torch.tensor([x_range[0], x_range[1], y_range[0], y_range[1], z_range[0], z_range[1]], dtype=torch.float64, device=dev)
    n_pts = n_grid ** 3
    ws = mlp_workspace(desc, _lib.PREC_BF16, n_pts, bool(sigma_only), False, dev)
    out = torch.empty(n_pts, 1 if sigma_only else 4, device=dev, dtype=torch.float32)
    ws_ptr, ws_bytes = _aligned(ws)
    check(lib.nerfb_mlp_fwd(C.byref(desc), _lib.PREC_BF16, _lib.IN_GRID, _req(params, "params"), 0, rg.data_ptr(), 0, n_pts,
                            int(n_grid), 1 if sigma_only else 0, 0, out.data_ptr(), ws_ptr, ws_bytes, _stream()), "nerfb_mlp_fwd[grid]")
    return out.view(n_grid, n_grid, n_grid) if sigma_only else out


def pack_vol(rgbsigma, neg_step):
    """extract_mesh.ipynb cell 7: (n, 4) [rgb sigma] -> uint32 (n_valid, 2) pairs (voxel index, packed rgba), index order."""
    n = rgbsigma.shape[0]
    dev = rgbsigma.device
    n_valid = torch.zeros(1, dtype=torch.int64, device=dev)
    if n == 0:
        return torch.empty(0, 2, dtype=torch.uint32, device=dev)
    ws = torch.empty(int(lib.nerfb_pack_vol_workspace_bytes(n)) + 1024, dtype=torch.uint8, device=dev)
    ws_ptr, ws_bytes = _aligned(ws)
    src = _req(rgbsigma, "rgbsigma")
    check(lib.nerfb_pack_vol(src, n, float(neg_step), 0, 0, n_valid.data_ptr(), ws_ptr, ws_bytes, _stream()), "nerfb_pack_vol[count]")
    k = int(n_valid.item())
    pairs = torch.empty(max(k, 1), 2, dtype=torch.uint32, device=dev)
    check(lib.nerfb_pack_vol(src, n, float(neg_step), pairs.data_ptr(), k, n_valid.data_ptr(), ws_ptr, ws_bytes, _stream()),
          "nerfb_pack_vol")
    return pairs[:k]


# ---- device guard ---------------------------------------------------------------------------------------------
# Kernels launch on the CUDA runtime's CURRENT device and on that device's current stream.  When the tensors of a call
# live on another GPU (evaluation on cuda:1 from a process whose current device is cuda:0), enter that device for the call.
def _on_tensor_device(fn):
    import functools

    @functools.wraps(fn)
    def guarded(*args, **kwargs):
        dev = None
        for a in args:
            if isinstance(a, torch.Tensor) and a.is_cuda:
                dev = a.device
                break
        if dev is None:
            for a in kwargs.values():
                if isinstance(a, torch.Tensor) and a.is_cuda:
                    dev = a.device
                    break
        if dev is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)
    return guarded


for _name in ("sample_coarse", "embed", "composite_fwd", "composite_bwd", "searchsorted", "sample_pdf", "resample",
              "adam_step", "mlp_fwd", "mlp_bwd", "render_fused", "triplane_relayout", "triplane_field", "triplane_field_bwd", "mip_march", "mip_march_bwd", "triplane_samples_bwd", "triplane_render",
              "finish_image", "sigma_grid", "pack_vol"):
    globals()[_name] = _on_tensor_device(globals()[_name])
del _name
