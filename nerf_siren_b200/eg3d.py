"""EG3D tri-plane renderer, host mirror of the reference's API:

    volumetric_rendering/renderer.py    generate_planes, sample_from_planes, ImportanceRenderer
    volumetric_rendering/ray_marcher.py MipRayMarcher2
    eg3d_training/triplane.py:144-167   OSGDecoder        (over networks_stylegan2.py:95-122 FullyConnectedLayer)

Same class / function names, argument meaning, state-dict keys (``net.0.weight`` ...) and output tuple as the
reference; all arithmetic runs in the hand-written kernels of csrc/triplane.cu (kernel 6).  Under ``torch.no_grad`` (the
reference's eval_eg3d.py:63-75) a render is ONE fused launch per batch element.  With grad enabled and planes / decoder
parameters that require grad (the reference's training step, system.py:111-120) the render is the same fused launch
that additionally keeps the samples (``nerfb_triplane_render_train``), and its backward is ``nerfb_mip_march_bwd`` (coarse
samples, then the depth-sorted union) followed by ``nerfb_triplane_field_bwd``; torch only does the sort / gather /
scatter of renderer.py:160-170.  ``run_model`` and ``MipRayMarcher2`` are differentiable on their own through the same
two backward kernels.  There is no CPU fallback and no torch-op fallback for the field or the marcher.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch
import torch.nn as nn

from . import ops


def generate_planes() -> torch.Tensor:
    """renderer.py:23-39 (kept for API parity; the kernel hard-codes these three projections, quirk included)."""
    return torch.tensor([[[1, 0, 0], [0, 1, 0], [0, 0, 1]],
                         [[1, 0, 0], [0, 0, 1], [0, 1, 0]],
                         [[0, 0, 1], [1, 0, 0], [0, 1, 0]]], dtype=torch.float32)


class FullyConnectedLayer(nn.Module):
    """Parameter container with the reference's init and gains (networks_stylegan2.py:95-122).  Only used as a
    sub-module of OSGDecoder here, so it has no forward of its own: the decoder runs fused inside kernel (6)."""

    def __init__(self, in_features, out_features, bias=True, activation="linear", lr_multiplier=1, bias_init=0):
        super().__init__()
        if activation != "linear" or not bias:
            raise NotImplementedError("only the linear+bias FullyConnectedLayer of OSGDecoder is on the hot path")
        self.in_features, self.out_features, self.activation = in_features, out_features, activation
        self.weight = nn.Parameter(torch.randn([out_features, in_features]) / lr_multiplier)
        self.bias = nn.Parameter(torch.full([out_features], np.float32(bias_init)))
        self.weight_gain = lr_multiplier / np.sqrt(in_features)
        self.bias_gain = lr_multiplier

    def extra_repr(self):
        return f"in_features={self.in_features:d}, out_features={self.out_features:d}, activation={self.activation:s}"


class OSGDecoder(nn.Module):
    """triplane.py:144-167.  ``forward(sampled_features (N, 3, M, 32), ray_directions)`` is served by the same fused
    decoder code as the renderer, on features that the caller has already gathered."""

    def __init__(self, n_features, options):
        super().__init__()
        if n_features != 32 or options.get("decoder_output_dim", 3) != 3:
            raise NotImplementedError("kernel (6) is built for 32 feature channels and 1 + 3 outputs")
        self.hidden_dim = 64
        self.lr_mul = float(options.get("decoder_lr_mul", 1))
        self.net = nn.Sequential(
            FullyConnectedLayer(n_features, self.hidden_dim, lr_multiplier=self.lr_mul),
            nn.Softplus(),
            FullyConnectedLayer(self.hidden_dim, 1 + options.get("decoder_output_dim", 3), lr_multiplier=self.lr_mul))

    def flat(self, detach: bool = True) -> torch.Tensor:
        """The 2372 raw parameters in the kernels' order.  ``detach=False`` keeps the autograd edge to the four tensors, so a
        gradient w.r.t. the flat vector is split back onto ``net.{0,2}.{weight,bias}`` by torch."""
        f = torch.cat([self.net[0].weight.reshape(-1), self.net[0].bias, self.net[2].weight.reshape(-1), self.net[2].bias])
        return (f.detach() if detach else f).float().contiguous()

    def wants_grad(self) -> bool:
        return any(p.requires_grad for p in self.parameters())

    def forward(self, sampled_features, ray_directions=None):
        raise NotImplementedError(
            "OSGDecoder is fused into the tri-plane kernel: call ImportanceRenderer.run_model(planes, decoder, coords, ...) "
            "(renderer.py:144-151) instead of gathering features first")


class _Planes:
    """Channels-last copy of one batch element's planes.  The copy is reused only for a PERSISTENT source -- the very same
    root tensor object (held by weak reference: a leaf / nn.Parameter such as ``LearnedPlanes.planes``) at the same
    ``_version``.  A generator backbone returns a fresh tensor per call, usually at the address the caching allocator just
    freed, so an address-based key would serve the previous call's planes: anything not provably the same object is
    re-laid out (one streaming kernel)."""

    def __init__(self):
        self.ref, self.version, self.index, self.cl = None, None, None, None

    def get(self, planes: torch.Tensor, b: int) -> torch.Tensor:
        import weakref
        root = planes._base if planes._base is not None else planes
        persistent = root.is_leaf and not (root.grad_fn is not None)
        if (persistent and self.ref is not None and self.ref() is root and self.version == root._version
                and self.index == (b, planes.data_ptr(), tuple(planes.shape))):
            return self.cl
        self.cl = ops.triplane_relayout(planes[b])
        self.ref, self.version = (weakref.ref(root), root._version) if persistent else (None, None)
        self.index = (b, planes.data_ptr(), tuple(planes.shape))
        return self.cl


def _training(planes, decoder) -> bool:
    """True when autograd expects this call to be differentiated (grad mode on and a plane / decoder leaf requires grad)."""
    return torch.is_grad_enabled() and (planes.requires_grad or decoder.wants_grad())


class _FieldFn(torch.autograd.Function):
    """run_model of one batch element (renderer.py:144-151, triplane.py:155-167) with its hand-written backward:
    forward nerfb_triplane_field, backward nerfb_triplane_field_bwd (recomputes the gather and the hidden layer per point,
    scatters into channels-last plane gradients).  ``planes`` is the reference layout (3, 32, H, W); coordinates get no
    gradient (the reference detaches the importance depths, renderer.py:214, and the coarse depths are constants)."""

    @staticmethod
    def forward(ctx, planes, dec_flat, lr_mul, coords, box_warp):
        cl = ops.triplane_relayout(planes)
        rgb, sigma = ops.triplane_field(cl, dec_flat, lr_mul, coords, box_warp)
        ctx.save_for_backward(cl, dec_flat, coords)
        ctx.lr_mul, ctx.box_warp = lr_mul, box_warp
        return rgb, sigma

    @staticmethod
    def backward(ctx, d_rgb, d_sigma):
        cl, dec_flat, coords = ctx.saved_tensors
        d_rgb = torch.zeros(coords.shape[0], 3, device=coords.device) if d_rgb is None else d_rgb.float().contiguous()
        d_sigma = torch.zeros(coords.shape[0], device=coords.device) if d_sigma is None else d_sigma.float().contiguous()
        d_cl, d_dec = ops.triplane_field_bwd(cl, dec_flat, ctx.lr_mul, coords, ctx.box_warp, d_rgb, d_sigma)
        return d_cl.permute(0, 3, 1, 2), d_dec, None, None, None


class _MarchFn(torch.autograd.Function):
    """MipRayMarcher2 (ray_marcher.py:25-57) on (rays, S) rows: forward nerfb_mip_march, backward nerfb_mip_march_bwd.
    Depths get no gradient (see _FieldFn)."""

    @staticmethod
    def forward(ctx, colors, densities, depths, white_back):
        rgb, depth, w = ops.mip_march(colors, densities, depths, white_back)
        ctx.save_for_backward(colors, densities, depths)
        ctx.white_back = white_back
        return rgb, depth, w

    @staticmethod
    def backward(ctx, d_rgb, d_depth, d_w):
        colors, densities, depths = ctx.saved_tensors
        c = lambda t: None if t is None else t.float().contiguous()
        d_c, d_s = ops.mip_march_bwd(colors, densities, depths, ctx.white_back, c(d_rgb), c(d_depth), c(d_w))
        return d_c, d_s, None, None


class _RenderFn(torch.autograd.Function):
    """ImportanceRenderer.forward of one batch element (renderer.py:94-142) as ONE autograd node: the forward is the fused
    render launch (nerfb_triplane_render_train: the six outputs plus every sample's depth / sigma / colour, 20 bytes per
    sample, the only thing kept for the backward); the backward is two launches: nerfb_triplane_samples_bwd (marcher backward
    on the coarse samples and on the depth-sorted union, renderer.py:160-170, per-sample cotangents and points) and
    nerfb_triplane_field_bwd.  Ray origins / directions / depths get no gradient (importance depths
    are detached in the reference, renderer.py:214)."""

    @staticmethod
    def forward(ctx, planes, dec_flat, lr_mul, o, d, S, Si, rs, re, start_arr, end_arr, disp, box_warp, wb, rand_coarse, u, seed):
        cl = ops.triplane_relayout(planes)
        *outs, samples = ops.triplane_render(cl, dec_flat, lr_mul, o, d, S, Si, rs, re, start_arr, end_arr, disp, box_warp, wb,
                                             rand_coarse, u, seed, return_samples=True)
        ctx.save_for_backward(cl, dec_flat, o, d, samples)
        ctx.cfg = (lr_mul, S, Si, box_warp, wb)
        return tuple(outs)

    @staticmethod
    def backward(ctx, g_rgb_c, g_depth_c, g_op_c, g_rgb_f, g_depth_f, g_op_f):
        cl, dec_flat, o, d, samples = ctx.saved_tensors
        lr_mul, S, Si, box_warp, wb = ctx.cfg
        c = lambda t: None if t is None else t.float().reshape(o.shape[0], -1).contiguous()
        coords, d_rgb, d_sig = ops.triplane_samples_bwd(samples, o, d, S, Si, wb, c(g_rgb_c), c(g_depth_c), c(g_op_c), c(g_rgb_f),
                                                        c(g_depth_f), c(g_op_f))
        d_cl, d_dec = ops.triplane_field_bwd(cl, dec_flat, lr_mul, coords, box_warp, d_rgb, d_sig)
        return (d_cl.permute(0, 3, 1, 2), d_dec) + (None,) * 15


def _march(colors, densities, depths, white_back):
    colors, densities, depths = colors.float().contiguous(), densities.float().contiguous(), depths.float().contiguous()
    if torch.is_grad_enabled() and (colors.requires_grad or densities.requires_grad):
        return _MarchFn.apply(colors, densities, depths, bool(white_back))
    return ops.mip_march(colors, densities, depths, bool(white_back))


class MipRayMarcher2(nn.Module):
    """ray_marcher.py:20-63."""

    def __init__(self):
        super().__init__()

    def run_forward(self, colors, densities, depths, rendering_options):
        if rendering_options["clamp_mode"] != "softplus":
            assert False, "MipRayMarcher only supports `clamp_mode`=`softplus`!"
        B, R, S, _ = depths.shape
        rgb, depth, w = _march(colors.reshape(B * R, S, 3), densities.reshape(B * R, S), depths.reshape(B * R, S),
                               bool(rendering_options.get("white_back", False)))
        return rgb.view(B, R, 3), depth.view(B, R, 1), w.view(B, R, S - 1, 1)

    def forward(self, colors, densities, depths, rendering_options):
        return self.run_forward(colors, densities, depths, rendering_options)


def get_ray_limits_box(rays_o, rays_d, box_side_length):
    """math_utils.py:48-99 restated with torch ops on the caller's device (host-side glue for ray_start == 'auto';
    slab test against the axis-aligned box, -1 / -2 for misses)."""
    o_shape = rays_o.shape
    o, d = rays_o.detach().reshape(-1, 3), rays_d.detach().reshape(-1, 3)
    half = box_side_length / 2
    inv = 1 / d
    lo = torch.where(inv < 0, torch.full_like(o, half), torch.full_like(o, -half))
    hi = torch.where(inv < 0, torch.full_like(o, -half), torch.full_like(o, half))
    t0, t1 = (lo - o) * inv, (hi - o) * inv
    valid = ~((t0[:, 0] > t1[:, 1]) | (t0[:, 1] > t1[:, 0]))
    tmin, tmax = torch.max(t0[:, 0], t0[:, 1]), torch.min(t1[:, 0], t1[:, 1])
    valid &= ~((tmin > t1[:, 2]) | (t0[:, 2] > tmax))
    tmin, tmax = torch.max(tmin, t0[:, 2]), torch.min(tmax, t1[:, 2])
    tmin = torch.where(valid, tmin, torch.full_like(tmin, -1.0))
    tmax = torch.where(valid, tmax, torch.full_like(tmax, -2.0))
    return tmin.reshape(*o_shape[:-1], 1), tmax.reshape(*o_shape[:-1], 1)


class ImportanceRenderer(nn.Module):
    """renderer.py:88-142.  ``forward`` returns the reference's 6-tuple
    (rgb_coarse, depth_coarse, weights_coarse.sum(2), rgb_final, depth_final, weights.sum(2))."""

    def __init__(self):
        super().__init__()
        self.ray_marcher = MipRayMarcher2()
        self.plane_axes = generate_planes()
        self._cache = _Planes()
        self.seed = 0

    def forward(self, planes, decoder, ray_origins, ray_directions, rendering_options, rand_coarse=None, u=None):
        opts = rendering_options
        if opts.get("clamp_mode", "softplus") != "softplus":
            assert False, "MipRayMarcher only supports `clamp_mode`=`softplus`!"
        if opts.get("density_noise", 0) > 0:
            raise NotImplementedError("density_noise > 0 is a training-time option; kernel (6) is the render path")
        n_imp = opts["depth_resolution_importance"]
        if n_imp <= 0:
            raise SystemExit(0)                     # the reference calls exit() here (renderer.py:138)
        B, R, _ = ray_origins.shape
        start_arr = end_arr = None
        rs, re = opts["ray_start"], opts["ray_end"]
        if isinstance(rs, str) and rs == re == "auto":
            start_arr, end_arr = get_ray_limits_box(ray_origins, ray_directions, box_side_length=opts["box_warp"])
            ok = end_arr > start_arr
            if torch.any(ok).item():                # renderer.py:101-103
                start_arr, end_arr = start_arr.clone(), end_arr.clone()
                start_arr[~ok] = start_arr[ok].min()
                end_arr[~ok] = start_arr[ok].max()
            rs = re = 0.0
        elif isinstance(rs, torch.Tensor):
            start_arr, end_arr, rs, re = rs, re, 0.0, 0.0
        outs = [[] for _ in range(6)]
        train = _training(planes, decoder)
        for b in range(B):                          # planes differ per batch element; rays of one element form one launch
            self.seed += 1
            if train:
                res = self._forward_train(planes[b], decoder, ray_origins[b].contiguous(), ray_directions[b].contiguous(), opts,
                                          float(rs), float(re), None if start_arr is None else start_arr[b].reshape(-1).contiguous(),
                                          None if end_arr is None else end_arr[b].reshape(-1).contiguous(),
                                          None if rand_coarse is None else rand_coarse[b].reshape(R, -1),
                                          None if u is None else u.reshape(B, R, -1)[b])
                for k in range(6):
                    outs[k].append(res[k])
                continue
            res = ops.triplane_render(
                self._cache.get(planes, b) if B == 1 else ops.triplane_relayout(planes[b]), decoder.flat(), decoder.lr_mul,
                ray_origins[b].contiguous(), ray_directions[b].contiguous(), opts["depth_resolution"], n_imp, float(rs), float(re),
                None if start_arr is None else start_arr[b].reshape(-1).contiguous(),
                None if end_arr is None else end_arr[b].reshape(-1).contiguous(),
                bool(opts.get("disparity_space_sampling", False)), float(opts["box_warp"]), bool(opts.get("white_back", False)),
                None if rand_coarse is None else rand_coarse[b].reshape(R, -1).contiguous(),
                None if u is None else u.reshape(B, R, -1)[b].contiguous(), self.seed)
            for k in range(6):
                outs[k].append(res[k])
        rgb_c, depth_c, op_c, rgb_f, depth_f, op_f = [torch.stack(o, 0) for o in outs]
        return rgb_c, depth_c.unsqueeze(-1), op_c.unsqueeze(-1), rgb_f, depth_f.unsqueeze(-1), op_f.unsqueeze(-1)

    def _forward_train(self, planes_b, decoder, o, d, opts, rs, re, start_arr, end_arr, rand_coarse, u):
        """renderer.py:94-142 for one batch element as one autograd node (_RenderFn): the same fused launch as under
        no_grad -- identical outputs -- which also keeps every sample for the backward."""
        return _RenderFn.apply(planes_b, decoder.flat(detach=False), decoder.lr_mul, o, d, int(opts["depth_resolution"]),
                               int(opts["depth_resolution_importance"]), rs, re, start_arr, end_arr,
                               bool(opts.get("disparity_space_sampling", False)), float(opts["box_warp"]),
                               bool(opts.get("white_back", False)),
                               None if rand_coarse is None else rand_coarse.float().contiguous(),
                               None if u is None else u.float().contiguous(), self.seed)

    def run_model(self, planes, decoder, sample_coordinates, sample_directions, options):
        """renderer.py:144-151 -> {'rgb': (N, M, 3), 'sigma': (N, M, 1)}."""
        if options.get("density_noise", 0) > 0:
            raise NotImplementedError("density_noise > 0 is not on the render path")
        N, M, _ = sample_coordinates.shape
        rgbs, sigmas = [], []
        if _training(planes, decoder):
            for b in range(N):
                rgb, sigma = _FieldFn.apply(planes[b], decoder.flat(detach=False), decoder.lr_mul,
                                            sample_coordinates[b].detach().float().contiguous(), float(options["box_warp"]))
                rgbs.append(rgb); sigmas.append(sigma)
            return {"rgb": torch.stack(rgbs, 0), "sigma": torch.stack(sigmas, 0).unsqueeze(-1)}
        for b in range(N):
            rgb, sigma = ops.triplane_field(ops.triplane_relayout(planes[b]), decoder.flat(), decoder.lr_mul,
                                            sample_coordinates[b].contiguous(), float(options["box_warp"]))
            rgbs.append(rgb); sigmas.append(sigma)
        return {"rgb": torch.stack(rgbs, 0), "sigma": torch.stack(sigmas, 0).unsqueeze(-1)}


# rendering_kwargs the reference hard-codes for its Unity scenes (eg3d_training/eg3d_renderer.py:30-36); only the fields the
# volume renderer reads
DEFAULT_RENDERING_KWARGS = {
    "depth_resolution": 64, "depth_resolution_importance": 64, "ray_start": 0.1, "ray_end": 10.0, "box_warp": 15.0,
    "white_back": False, "disparity_space_sampling": False, "clamp_mode": "softplus", "decoder_lr_mul": 1.0,
}


class LearnedPlanes(nn.Module):
    """Stand-in for the plane source: a (1, 96, H, W) feature image held as a parameter (so it moves with ``.to()`` and
    lands in checkpoints).  It trains through the renderer's differentiable schedule (plane gradients come from
    nerfb_triplane_field_bwd).  The reference produces the planes
    with a StyleGAN2 backbone (``TriPlaneGenerator.backbone.synthesis``, eg3d_training/triplane.py:61), which is outside the
    render_rays path (out of scope); any module / callable returning (B, 96, H, W) can take its place."""

    def __init__(self, resolution=256, channels=96):
        super().__init__()
        self.planes = nn.Parameter(torch.randn(1, channels, resolution, resolution))

    def forward(self, *_a, **_k):
        return self.planes


class EG3D_Renderer(nn.Module):
    """eg3d_training/eg3d_renderer.py:6-67: the object ``NeRFSystem_eg3d`` renders with (system.py:22,75) and
    ``extract_color_mesh_eg3d.py:171-190`` samples.  ``render`` returns the reference's six keys with the batch axis
    squeezed; ``sample`` returns run_model's {'rgb', 'sigma'}.  ``plane_source``: module / callable -> (1, 96, H, W)."""

    def __init__(self, plane_source=None, rendering_kwargs=None, decoder=None):
        super().__init__()
        self.rendering_kwargs = dict(DEFAULT_RENDERING_KWARGS if rendering_kwargs is None else rendering_kwargs)
        self.plane_source = LearnedPlanes() if plane_source is None else plane_source
        self.decoder = decoder if decoder is not None else OSGDecoder(
            32, {"decoder_lr_mul": self.rendering_kwargs.get("decoder_lr_mul", 1.0), "decoder_output_dim": 3})
        self.renderer = ImportanceRenderer()

    def _planes(self, conditioning_params):
        planes = self.plane_source(conditioning_params) if callable(self.plane_source) else self.plane_source
        return planes.view(len(planes), 3, 32, planes.shape[-2], planes.shape[-1])       # triplane.py:65

    def render(self, conditioning_params, ray_origins, ray_directions, **kw):
        out = self.renderer(self._planes(conditioning_params), self.decoder, ray_origins.unsqueeze(0),
                            ray_directions.unsqueeze(0), self.rendering_kwargs, **kw)
        keys = ("rgb_coarse", "depth_coarse", "opacity_coarse", "rgb_fine", "depth_fine", "opacity_fine")
        return {k: v.squeeze(0) for k, v in zip(keys, out)}

    def sample(self, coordinates, directions):
        return self.renderer.run_model(self._planes(-1), self.decoder, coordinates, directions, self.rendering_kwargs)
