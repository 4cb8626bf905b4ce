"""Host glue between nn.Modules carrying the reference's parameters and the C-ABI MLP entry points."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib, config, ops
from ._lib import MlpDesc


def flat_params(module: torch.nn.Module) -> torch.Tensor:
    """One contiguous fp32 blob in state_dict order; torch.cat keeps autograd routing grads to the params."""
    sl = getattr(module, "_nerfb_flat_slice", None)
    if sl is not None:                       # parameters already live in one flat buffer (dp.FlatParameters)
        flat = sl.get(module)
        if flat is not None:
            return flat
    ps = list(module.parameters())
    for p in ps:
        if not p.is_cuda:
            raise RuntimeError("model parameters must live on a CUDA device (no CPU path): call .cuda() first")
        if p.dtype != torch.float32:
            raise ValueError("model parameters must be float32")
    return torch.cat([p.reshape(-1) for p in ps])


def pick_precision(desc: MlpDesc, need_grad: bool) -> int:
    want = config.get_precision()
    if want == "fp32":
        return _lib.PREC_FP32
    ok = ops.mlp_supported(desc, _lib.PREC_BF16, need_grad)
    if want == "bf16" and not ok:
        raise RuntimeError("precision 'bf16' requested but this model shape (or its backward) is outside the "
                           "tcgen05 kernels")
    if not ok:
        _warn_fp32_once(desc, need_grad)
    return _lib.PREC_BF16 if ok else _lib.PREC_FP32


_warned = set()


def _warn_fp32_once(desc: MlpDesc, need_grad: bool) -> None:
    """'auto' silently taking the fp32 CUDA-core arm is a ~30x cliff (W != 256, CLS > 8, D > 9, ...): say so once per shape."""
    key = (desc.kind, desc.D, desc.W, desc.n_cls, bool(need_grad))
    if key in _warned:
        return
    _warned.add(key)
    import warnings
    warnings.warn(f"nerf_siren_b200: model shape (kind={desc.kind}, D={desc.D}, W={desc.W}, CLS={desc.n_cls}, "
                  f"grad={bool(need_grad)}) is outside the tcgen05 tensor-core kernels; precision 'auto' falls back to the "
                  "fp32 CUDA-core arm, which is roughly 30x slower. Supported: W=256, D<=9, CLS<=8.", RuntimeWarning,
                  stacklevel=3)


def max_points(precision: int) -> int:
    return config.MAX_POINTS_FP32 if precision == _lib.PREC_FP32 else config.MAX_POINTS_BF16


class MLPFunction(torch.autograd.Function):
    """out = MLP(x) for an already-embedded x (NeRF.forward called directly, extract_color_mesh.py:136-139)."""

    @staticmethod
    def forward(ctx, flat, x, film, desc, precision, sigma_only):
        need = bool(ctx.needs_input_grad[0] or ctx.needs_input_grad[1]) and not sigma_only
        if ctx.needs_input_grad[2]:
            raise RuntimeError("gradients w.r.t. the FiLM frequencies / phase shifts are not implemented (the reference's "
                               "mapping network is commented out, models/nerf.py:185): detach them")
        if ctx.needs_input_grad[1] and desc.kind == _lib.KIND_SIREN:
            raise RuntimeError("gradients w.r.t. the SIREN input (points / directions) are not implemented")
        if ctx.needs_input_grad[1] and precision != _lib.PREC_FP32:
            raise RuntimeError("gradients w.r.t. the embedded input are only implemented by the fp32 arm")
        ws = ops.mlp_workspace(desc, precision, x.shape[0], sigma_only, need, x.device)
        out = ops.mlp_fwd(desc, precision, _lib.IN_EMBEDDED, flat, film, x, None, x.shape[0], 1, sigma_only, need, ws)
        if need:
            ctx.save_for_backward(flat, x, out, ws, film if film is not None else torch.empty(0, device=x.device))
            ctx.desc, ctx.precision, ctx.has_film = desc, precision, film is not None
        ctx.sigma_only = sigma_only
        return out

    @staticmethod
    def backward(ctx, d_out):
        if ctx.sigma_only:
            raise RuntimeError("sigma_only outputs are inference-only (the reference uses them under test_time)")
        flat, x, out, ws, film = ctx.saved_tensors
        grad, d_in = ops.mlp_bwd(ctx.desc, ctx.precision, _lib.IN_EMBEDDED, flat, film if ctx.has_film else None, x,
                                 None, x.shape[0], 1, out, d_out.contiguous(), ws, want_d_in=ctx.needs_input_grad[1])
        return grad, d_in, None, None, None, None


def run_embedded(module, desc: MlpDesc, x: torch.Tensor, sigma_only: bool, film: Optional[torch.Tensor] = None):
    if x.dim() != 2:
        raise ValueError(f"expected (B, C) input, got {tuple(x.shape)}")
    if not x.is_cuda:
        raise RuntimeError("input must be a CUDA tensor (nerf_siren_b200 has no CPU path)")
    x = x.contiguous().float()
    need_grad = torch.is_grad_enabled() and not sigma_only and (
        x.requires_grad or any(p.requires_grad for p in module.parameters()))
    precision = pick_precision(desc, need_grad)
    flat = flat_params(module)
    cap = max_points(precision)
    if x.shape[0] <= cap:
        return MLPFunction.apply(flat, x, film, desc, precision, sigma_only)
    return torch.cat([MLPFunction.apply(flat, x[i:i + cap], film, desc, precision, sigma_only)
                      for i in range(0, x.shape[0], cap)], 0)
