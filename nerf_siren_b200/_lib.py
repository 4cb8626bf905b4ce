"""ctypes binding of libnerf_b200.so (C ABI declared in include/nerf_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a only).  There is no CPU or
PyTorch fallback: if the shared object is missing this module raises at import time, and every call
checks the returned status and raises ``RuntimeError`` with the library's own message.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnerf_b200.so")


class MlpDesc(C.Structure):
    """Mirror of ``nerfb_mlp_desc`` (include/nerf_b200.h)."""
    _fields_ = [("kind", C.c_int32), ("D", C.c_int32), ("W", C.c_int32), ("n_freq_xyz", C.c_int32),
                ("n_freq_dir", C.c_int32), ("skip_mask", C.c_int32), ("n_cls", C.c_int32),
                ("logscale", C.c_int32), ("box", C.c_float), ("in_xyz", C.c_int32), ("in_dir", C.c_int32)]


class RenderCfg(C.Structure):
    """Mirror of ``nerfb_render_cfg``."""
    _fields_ = [("precision", C.c_int32), ("n_samples", C.c_int32), ("n_importance", C.c_int32), ("use_disp", C.c_int32),
                ("white_back", C.c_int32), ("test_time", C.c_int32), ("perturb", C.c_float), ("noise_std", C.c_float)]


class RenderIO(C.Structure):
    """Mirror of ``nerfb_render_io`` (device pointers, 0 = NULL)."""
    _fields_ = [(k, C.c_void_p) for k in ("rays", "perturb_rand", "noise_coarse", "u", "noise_fine", "rgb_coarse",
                                          "depth_coarse", "opacity_coarse", "cls_coarse", "rgb_fine", "depth_fine",
                                          "opacity_fine", "cls_fine")]


KIND_NERF, KIND_SIREN = 0, 1
PREC_FP32, PREC_BF16 = 0, 1
IN_RAYS, IN_EMBEDDED, IN_GRID = 0, 1, 2
E_UNSUPPORTED = -2

_P, _I, _L, _F = C.c_void_p, C.c_int, C.c_int64, C.c_float
_D = C.POINTER(MlpDesc)

# name -> (restype, argtypes); the authoritative list of exported symbols (tests check it against the header)
SIGNATURES = {
    "nerfb_version": (_I, []),
    "nerfb_arch": (_I, []),
    "nerfb_error_string": (C.c_char_p, [_I]),
    "nerfb_param_count": (_L, [_D]),
    "nerfb_out_channels": (_I, [_D]),
    "nerfb_launch_count": (_L, []),
    "nerfb_sample_coarse": (_I, [_P, _P, _F, _I, _L, _I, _P, _P]),
    "nerfb_embed": (_I, [_P, _L, _I, _I, _I, _P, _P]),
    "nerfb_mlp_workspace_bytes": (_L, [_D, _I, _L, _I, _I]),
    "nerfb_mlp_fwd": (_I, [_D, _I, _I, _P, _P, _P, _P, _L, _I, _I, _I, _P, _P, _L, _P]),
    "nerfb_mlp_bwd": (_I, [_D, _I, _I, _P, _P, _P, _P, _L, _I, _P, _P, _P, _P, _P, _L, _P]),
    "nerfb_render_fused_workspace_bytes": (_L, [_D, _D, C.POINTER(RenderCfg), _L]),
    "nerfb_render_fused_fwd": (_I, [_D, _D, C.POINTER(RenderCfg), _P, _P, _P, _P, C.POINTER(RenderIO), _L, _P, _L, _P]),
    "nerfb_composite_fwd": (_I, [_P, _I, _P, _P, _P, _F, _I, _L, _I, _P, _P, _P, _P, _P, _P]),
    "nerfb_composite_bwd": (_I, [_P, _I, _P, _P, _P, _F, _I, _L, _I, _P, _P, _P, _P, _P, _P, _P]),
    "nerfb_searchsorted": (_I, [_P, _L, _L, _P, _L, _L, _I, _P, _P]),
    "nerfb_searchsorted_typed": (_I, [_P, _L, _L, _P, _L, _L, _I, _I, _P, _P]),
    "nerfb_sample_pdf": (_I, [_P, _P, _P, _I, _F, _L, _I, _I, _P, _P, _P, _P]),
    "nerfb_resample": (_I, [_P, _P, _P, _I, _L, _I, _I, _P, _P]),
    "nerfb_triplane_relayout": (_I, [_P, _I, _I, _I, _P, _P]),
    "nerfb_triplane_field": (_I, [_P, _I, _I, _P, _F, _P, _L, _F, _P, _P, _P]),
    "nerfb_mip_march": (_I, [_P, _P, _P, _L, _I, _I, _P, _P, _P, _P, _P]),
    "nerfb_triplane_field_bwd": (_I, [_P, _I, _I, _P, _F, _P, _L, _F, _P, _P, _P, _P, _P]),
    "nerfb_mip_march_bwd": (_I, [_P, _P, _P, _L, _I, _I, _P, _P, _P, _P, _P, _P]),
    "nerfb_triplane_render": (_I, [_P, _I, _I, _P, _F, _P, _P, _L, _I, _I, C.c_double, C.c_double, _P, _P, _I, _F, _I,
                                   _P, _P, C.c_uint64, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "nerfb_triplane_samples_bwd": (_I, [_P, _P, _P, _L, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "nerfb_triplane_render_train": (_I, [_P, _I, _I, _P, _F, _P, _P, _L, _I, _I, C.c_double, C.c_double, _P, _P, _I, _F, _I,
                                         _P, _P, C.c_uint64, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "nerfb_generate_rays": (_I, [_P, _I, _I, C.c_double, _F, _F, _I, _P, _P]),
    "nerfb_finish_image": (_I, [_P, _P, _P, _I, _I, _P, _P, _P, _P]),
    "nerfb_pack_vol_workspace_bytes": (_L, [_L]),
    "nerfb_pack_vol": (_I, [_P, _L, _F, _P, _L, _P, _P, _L, _P]),
    "nerfb_pack_vol_static_max_voxels": (_L, []),
    "nerfb_pack_vol_static": (_I, [_P, _L, _F, _P, _L, _P, _P]),
    "nerfb_adam_step": (_I, [_P, _P, _P, _P, _L, C.c_double, C.c_double, C.c_double, C.c_double, _L, _F, _P]),
}

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
        "`python -c 'import __graft_entry__ as g; g.build()'` from the repo root (needs nvcc). "
        "nerf_siren_b200 has no CPU / PyTorch fallback.")

lib = C.CDLL(LIB_PATH)
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)          # AttributeError here == the library does not export the ABI
    _fn.restype = _res
    _fn.argtypes = _args


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib.nerfb_error_string(int(status)).decode()
        raise RuntimeError(f"{what} failed with status {status}: {msg}")
