"""Drop-in for the reference's ``models/rendering.py``: ``render_rays`` / ``sample_pdf`` (+ the ``_3d`` names
``system.py:5`` and ``eval.py:9`` import) with the reference's signatures, result-dict keys and RNG draw
order, running on the sm_100a kernels of libnerf_b200.so.

    reference                                      here
    models/rendering.py:206-222  z_vals            kernel (2) nerfb_sample_coarse
    :129-159 embed / cat / model chunk loop        kernel (1) nerfb_mlp_fwd / nerfb_mlp_bwd (fused encoding + MLP)
    :161-190 alpha compositing (+ :379 classes)    kernel (4) nerfb_composite_fwd / _bwd
    :241-247 sample_pdf + cat + sort               kernel (3) nerfb_resample

Two host paths, both CUDA-only:
  * fused   -- models are this package's NeRF / NeRF_3D / SirenField and embeddings its Embedding: one
               autograd node per ray chunk, xyz / encodings / per-layer activations never visit HBM in fp32;
  * generic -- any other callables ``model(x, sigma_only=...)`` / ``embedding(x)``: the reference's chunk loop
               around the black boxes, with kernels (2)(3)(4) for everything else.
"""
from __future__ import annotations

import torch

from .. import _field, _lib, config, ops
from .nerf import Embedding, NeRF, SirenField

__all__ = ["render_rays", "render_rays_3d", "render_rays_3d_conv", "sample_pdf"]


def sample_pdf(bins, weights, N_importance, det=False, eps=1e-5):
    """models/rendering.py:22-67.  bins (N, n+1), weights (N, n) -> samples (N, N_importance).
    Like the reference, the result carries no gradient path that callers rely on (they .detach() it, :244)."""
    bins, weights = bins.contiguous().float(), weights.detach().contiguous().float()
    return ops.sample_pdf(bins, weights, int(N_importance), det=bool(det), eps=float(eps))


# ------------------------------------------------------------------------------------------------------
# autograd nodes
# ------------------------------------------------------------------------------------------------------
class _CompositeFn(torch.autograd.Function):
    """(raw (N,S,C), z) -> weights, rgb, depth, opacity, extra with the fused backward of kernel (4)."""

    @staticmethod
    def forward(ctx, raw, z, rays, noise, noise_std, white_back):
        ctx.set_materialize_grads(False)
        raw = raw.contiguous()
        w, rgb, depth, opacity, extra = ops.composite_fwd(raw, z, rays, noise, noise_std, white_back)
        ctx.save_for_backward(raw, z, rays, noise if noise is not None else torch.empty(0, device=raw.device))
        ctx.cfg = (noise is not None, noise_std, white_back)
        outs = (w, rgb, depth, opacity, extra)
        return tuple(o if o is not None else torch.empty(0, device=raw.device) for o in outs)

    @staticmethod
    def backward(ctx, gw, grgb, gdepth, gop, gextra):
        raw, z, rays, noise = ctx.saved_tensors
        has_noise, noise_std, white_back = ctx.cfg
        c = raw.shape[-1]

        def opt(g, used):
            return g.contiguous() if (used and g is not None) else None
        d_raw = ops.composite_bwd(raw, z, rays, noise if has_noise else None, noise_std, white_back,
                                  d_rgb=opt(grgb, c != 1), d_depth=opt(gdepth, c != 1), d_opacity=opt(gop, True),
                                  d_weights=opt(gw, True), d_extra=opt(gextra, c > 4))
        return d_raw, None, None, None, None, None


class _Cfg:
    __slots__ = ("S", "Si", "use_disp", "perturb", "noise_std", "white_back", "test_time", "desc_c", "desc_f",
                 "precision", "film_c", "film_f")


class _RenderFn(torch.autograd.Function):
    """One ray chunk of the fused coarse(+fine) render.  Differentiable inputs: the two flat parameter blobs."""

    @staticmethod
    def forward(ctx, flat_c, flat_f, rays, cfg, perturb_rand, noise_c, u, noise_f):
        ctx.set_materialize_grads(False)
        n, S, Si = rays.shape[0], cfg.S, cfg.Si
        # (torch.is_grad_enabled() is always False inside Function.forward: ask the node instead)
        need_c = bool(ctx.needs_input_grad[0]) and not cfg.test_time
        need_f = bool(ctx.needs_input_grad[1]) and flat_f is not None and Si > 0
        nz_c = noise_c if cfg.noise_std != 0 else None
        nz_f = noise_f if cfg.noise_std != 0 else None

        z = ops.sample_coarse(rays, S, cfg.use_disp, cfg.perturb, perturb_rand)
        ws_c = ops.mlp_workspace(cfg.desc_c, cfg.precision, n * S, cfg.test_time, need_c, rays.device)
        raw_c = ops.mlp_fwd(cfg.desc_c, cfg.precision, _lib.IN_RAYS, flat_c, cfg.film_c, rays, z, n, S, cfg.test_time,
                            need_c, ws_c).view(n, S, -1)
        w_c, rgb_c, depth_c, op_c, cls_c = ops.composite_fwd(raw_c, z, rays, nz_c, cfg.noise_std, cfg.white_back)
        rgb_f = depth_f = op_f = cls_f = z_f = raw_f = ws_f = None
        if Si > 0:
            z_f = ops.resample(z, w_c, Si, det=(cfg.perturb == 0), u=u)                  # rendering.py:241-247
            ws_f = ops.mlp_workspace(cfg.desc_f, cfg.precision, n * (S + Si), False, need_f, rays.device)
            raw_f = ops.mlp_fwd(cfg.desc_f, cfg.precision, _lib.IN_RAYS, flat_f, cfg.film_f, rays, z_f, n, S + Si, False,
                                need_f, ws_f).view(n, S + Si, -1)
            _, rgb_f, depth_f, op_f, cls_f = ops.composite_fwd(raw_f, z_f, rays, nz_f, cfg.noise_std, cfg.white_back)
        ctx.cfg, ctx.need_c, ctx.need_f = cfg, need_c, need_f
        if need_c or need_f:
            e = torch.empty(0, device=rays.device)
            ctx.save_for_backward(flat_c, flat_f if flat_f is not None else e, rays, z, raw_c, ws_c if need_c else e,
                                  nz_c if nz_c is not None else e, z_f if z_f is not None else e,
                                  raw_f if raw_f is not None else e, ws_f if need_f else e,
                                  nz_f if nz_f is not None else e)
        outs = (rgb_c, depth_c, op_c, cls_c, rgb_f, depth_f, op_f, cls_f)
        ctx.present = tuple(o is not None for o in outs)
        e = torch.empty(0, device=rays.device)
        return tuple(o if o is not None else e for o in outs)

    @staticmethod
    def backward(ctx, g_rgb_c, g_depth_c, g_op_c, g_cls_c, g_rgb_f, g_depth_f, g_op_f, g_cls_f):
        cfg = ctx.cfg
        flat_c, flat_f, rays, z, raw_c, ws_c, nz_c, z_f, raw_f, ws_f, nz_f = ctx.saved_tensors
        n, S, Si = rays.shape[0], cfg.S, cfg.Si
        p = ctx.present

        def opt(g, present):
            return g.contiguous() if (present and g is not None) else None

        def arm(flat, film, desc, raw, zz, noise, ws, s, grads, present):
            c = raw.shape[-1]
            d_raw = ops.composite_bwd(raw, zz, rays, noise if noise.numel() else None, cfg.noise_std, cfg.white_back,
                                      d_rgb=opt(grads[0], present[0]), d_depth=opt(grads[1], present[1]),
                                      d_opacity=opt(grads[2], present[2]),
                                      d_extra=opt(grads[3], present[3] and c > 4))
            g, _ = ops.mlp_bwd(desc, cfg.precision, _lib.IN_RAYS, flat, film, rays, zz, n, s, raw.view(n * s, c),
                               d_raw.view(n * s, c), ws)
            return g

        g_c = g_f = None
        if ctx.need_f:
            g_f = arm(flat_f, cfg.film_f, cfg.desc_f, raw_f, z_f, nz_f, ws_f, S + Si,
                      (g_rgb_f, g_depth_f, g_op_f, g_cls_f), p[4:8])
        if ctx.need_c:
            g_c = arm(flat_c, cfg.film_c, cfg.desc_c, raw_c, z, nz_c, ws_c, S,
                      (g_rgb_c, g_depth_c, g_op_c, g_cls_c), p[0:4])
        return g_c, g_f, None, None, None, None, None, None


# ------------------------------------------------------------------------------------------------------
def _fusable(model, emb_xyz, emb_dir) -> bool:
    if not (isinstance(emb_xyz, Embedding) and isinstance(emb_dir, Embedding)):
        return False
    if emb_xyz.in_channels != 3 or emb_dir.in_channels != 3 or emb_xyz.logscale != emb_dir.logscale:
        return False
    if isinstance(model, SirenField):
        return emb_xyz.N_freqs == 0 and emb_dir.N_freqs == 0
    if isinstance(model, NeRF):
        return model.in_channels_xyz == emb_xyz.out_channels and model.in_channels_dir == emb_dir.out_channels
    return False


def _check_rays(rays):
    if rays.dim() != 2 or rays.shape[1] != 8:
        raise ValueError(f"rays must be (N_rays, 8) [o, d, near, far], got {tuple(rays.shape)}")
    if not rays.is_cuda:
        raise RuntimeError("rays must be a CUDA tensor: nerf_siren_b200 has no CPU path")
    return rays.contiguous().float()


def _render_fused(models, embeddings, rays, S, use_disp, perturb, noise_std, Si, white_back, test_time, want_cls):
    emb_xyz, emb_dir = embeddings[0], embeddings[1]
    mc = models[0]
    mf = models[1] if Si > 0 else None
    cfg = _Cfg()
    cfg.S, cfg.Si, cfg.use_disp, cfg.perturb = int(S), int(Si), bool(use_disp), float(perturb)
    cfg.noise_std, cfg.white_back, cfg.test_time = float(noise_std), bool(white_back), bool(test_time)
    cfg.desc_c = mc._desc(False, emb_xyz.N_freqs, emb_dir.N_freqs, emb_xyz.logscale)
    cfg.desc_f = mf._desc(False, emb_xyz.N_freqs, emb_dir.N_freqs, emb_xyz.logscale) if mf is not None else None
    cfg.film_c = mc._film()
    cfg.film_f = mf._film() if mf is not None else None
    need_grad = torch.is_grad_enabled() and not test_time and any(
        p.requires_grad for m in (mc, mf) if m is not None for p in m.parameters())
    pc = _field.pick_precision(cfg.desc_c, need_grad)
    pf = _field.pick_precision(cfg.desc_f, need_grad) if mf is not None else pc
    cfg.precision = pc if pc == pf else _lib.PREC_FP32
    flat_c = _field.flat_params(mc)
    flat_f = _field.flat_params(mf) if mf is not None else None

    n = rays.shape[0]
    dev = rays.device
    # RNG stream parity with the reference: same draws, same shapes, same order, even when unused (:170)
    perturb_rand = torch.rand(n, S, device=dev) if perturb > 0 else None
    noise_c = torch.randn(n, S, device=dev)
    u = torch.rand(n, Si, device=dev) if (Si > 0 and perturb > 0) else None
    noise_f = torch.randn(n, S + Si, device=dev) if Si > 0 else None

    rays_per_call = max(1, _field.max_points(cfg.precision) // (S + Si))
    pieces = []
    e = torch.empty(0, device=dev)
    for i in range(0, n, rays_per_call):
        sl = slice(i, i + rays_per_call)
        if not need_grad and config.RENDER_FUSED:
            # inference: the whole chain behind one library call (nerfb_render_fused_fwd), compositing in the MLP epilogue
            # where the sample counts allow it
            piece = ops.render_fused(cfg.desc_c, cfg.desc_f, cfg.precision, flat_c.detach(), cfg.film_c,
                                     None if flat_f is None else flat_f.detach(), cfg.film_f, rays[sl], S, Si, cfg.use_disp,
                                     cfg.perturb, cfg.noise_std, cfg.white_back, cfg.test_time,
                                     perturb_rand[sl] if perturb_rand is not None else None,
                                     noise_c[sl] if cfg.noise_std != 0 else None, u[sl] if u is not None else None,
                                     noise_f[sl] if (noise_f is not None and cfg.noise_std != 0) else None)
            pieces.append(tuple(o if o is not None else e for o in piece))
            continue
        pieces.append(_RenderFn.apply(flat_c, flat_f, rays[sl], cfg,
                                      perturb_rand[sl] if perturb_rand is not None else None, noise_c[sl],
                                      u[sl] if u is not None else None, noise_f[sl] if noise_f is not None else None))
    outs = [torch.cat([p[k] for p in pieces], 0) if len(pieces) > 1 else pieces[0][k] for k in range(8)]
    rgb_c, depth_c, op_c, cls_c, rgb_f, depth_f, op_f, cls_f = outs
    has_cls = cls_c.numel() > 0 or cls_f.numel() > 0
    if test_time:
        result = {"opacity_coarse": op_c}
    else:
        result = {"rgb_coarse": rgb_c, "depth_coarse": depth_c}
        if has_cls and want_cls:
            result["cls_coarse"] = cls_c
        result["opacity_coarse"] = op_c
    if Si > 0:
        result["rgb_fine"], result["depth_fine"] = rgb_f, depth_f
        if has_cls and want_cls:
            result["cls_fine"] = cls_f
        result["opacity_fine"] = op_f
    return result


def _render_generic(models, embeddings, rays, S, use_disp, perturb, noise_std, Si, chunk, white_back, test_time,
                    want_cls):
    """Black-box models: the reference's chunk loop (rendering.py:129-152) around kernels (2)(3)(4)."""
    emb_xyz, emb_dir = embeddings[0], embeddings[1]
    n, dev = rays.shape[0], rays.device
    rays_o, rays_d = rays[:, 0:3], rays[:, 3:6]
    dir_emb = emb_dir(rays_d.contiguous())

    def shade(model, z, noise, sigma_only):
        s = z.shape[1]
        xyz = (rays_o.unsqueeze(1) + rays_d.unsqueeze(1) * z.unsqueeze(2)).reshape(-1, 3)
        rep = None if sigma_only else torch.repeat_interleave(dir_emb, repeats=s, dim=0)
        outs = []
        for i in range(0, xyz.shape[0], chunk):
            e = emb_xyz(xyz[i:i + chunk].contiguous())
            if not sigma_only:
                e = torch.cat([e, rep[i:i + chunk]], 1)
            outs.append(model(e, sigma_only=sigma_only))
        raw = torch.cat(outs, 0).view(n, s, -1).float()
        return _CompositeFn.apply(raw, z, rays, noise if noise_std != 0 else None, float(noise_std), bool(white_back))

    perturb_rand = torch.rand(n, S, device=dev) if perturb > 0 else None
    z = ops.sample_coarse(rays, S, use_disp, perturb, perturb_rand)
    noise_c = torch.randn(n, S, device=dev)
    w_c, rgb_c, depth_c, op_c, cls_c = shade(models[0], z, noise_c, test_time)
    if test_time:
        result = {"opacity_coarse": op_c}
    else:
        result = {"rgb_coarse": rgb_c, "depth_coarse": depth_c}
        if want_cls and cls_c.numel():
            result["cls_coarse"] = cls_c
        result["opacity_coarse"] = op_c
    if Si > 0:
        u = torch.rand(n, Si, device=dev) if perturb > 0 else None
        z_f = ops.resample(z, w_c.detach().contiguous(), Si, det=(perturb == 0), u=u)
        noise_f = torch.randn(n, S + Si, device=dev)
        _, rgb_f, depth_f, op_f, cls_f = shade(models[1], z_f, noise_f, False)
        result["rgb_fine"], result["depth_fine"] = rgb_f, depth_f
        if want_cls and cls_f.numel():
            result["cls_fine"] = cls_f
        result["opacity_fine"] = op_f
    return result


def _render(models, embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance, chunk, white_back,
            test_time, want_cls):
    rays = _check_rays(rays)
    if N_importance > 0 and len(models) < 2:
        raise ValueError("N_importance > 0 needs a fine model: models = [coarse, fine]")
    if N_importance > 0 and N_samples < 3:
        raise ValueError("hierarchical sampling needs at least 3 coarse samples")
    if rays.shape[0] == 0:                       # empty batch: same keys, empty tensors, no launch
        e1, e3 = rays.new_zeros(0), rays.new_zeros(0, 3)
        out = {"opacity_coarse": e1} if test_time else {"rgb_coarse": e3, "depth_coarse": e1, "opacity_coarse": e1}
        if N_importance > 0:
            out.update({"rgb_fine": e3, "depth_fine": e1, "opacity_fine": e1})
        return out
    fused = _fusable(models[0], embeddings[0], embeddings[1]) and (
        N_importance == 0 or (_fusable(models[1], embeddings[0], embeddings[1]) and type(models[1]) is type(models[0])))
    if fused:
        return _render_fused(models, embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance,
                             white_back, test_time, want_cls)
    return _render_generic(models, embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance, chunk,
                           white_back, test_time, want_cls)


def render_rays(models, embeddings, rays, N_samples=64, use_disp=False, perturb=0, noise_std=1, N_importance=0,
                chunk=1024 * 32, white_back=False, test_time=False, _cls_num=6, network=None):
    """models/rendering.py:70-262.  Result keys: 'opacity_coarse' always; 'rgb_coarse', 'depth_coarse' unless
    test_time; 'rgb_fine', 'depth_fine', 'opacity_fine' iff N_importance > 0.  ``_cls_num`` / ``network`` are
    accepted and ignored, as in the reference.  ``chunk`` only matters for black-box models."""
    return _render(models, embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance, chunk, white_back,
                   test_time, want_cls=False)


def render_rays_3d(models, embeddings, rays, N_samples=64, use_disp=False, perturb=0, noise_std=1, N_importance=0,
                   chunk=1024 * 32, white_back=False, test_time=False, _cls_num=6, no_grad_on_nerf=True, network=None):
    """The class-rendering entry (models/rendering.py:265-472): adds 'cls_coarse' / 'cls_fine' = sum_i w_i cls_i
    (:379).  The per-sample class scores are the NeRF_3D head's channels 4: (BASELINE config 4); the reference's
    PointNet-over-thresholded-points branch (:350-377) is outside the hot path and not provided (``network`` is
    accepted and ignored).  ``no_grad_on_nerf=True`` -- the reference's default -- runs the NeRF models under
    ``torch.no_grad()`` exactly as :306-308 does, so no gradient reaches them (in the reference only the PointNet
    trains in that mode); pass ``no_grad_on_nerf=False`` to train the NeRF_3D trunk and class head."""
    if not hasattr(models[0], "CLS") or models[0].CLS == 0:
        raise NotImplementedError("render_rays_3d here composites the NeRF_3D class head; the reference's PointNet "
                                  "branch (models/rendering.py:350-377) is out of scope")
    if no_grad_on_nerf:
        with torch.no_grad():
            return _render(models[:2], embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance, chunk,
                           white_back, test_time, want_cls=True)
    return _render(models[:2], embeddings, rays, N_samples, use_disp, perturb, noise_std, N_importance, chunk,
                   white_back, test_time, want_cls=True)


def render_rays_3d_conv(*args, **kwargs):
    """models/rendering.py:475-732 differs from render_rays_3d only in its MinkowskiEngine branch (commented out
    upstream, :609-643) and a log_softmax on the class map (:651-652)."""
    out = render_rays_3d(*args, **kwargs)
    for k in ("cls_coarse", "cls_fine"):
        if k in out:
            out[k] = torch.log_softmax(out[k], dim=-1)
    return out
