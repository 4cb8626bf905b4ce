"""nerf_siren_b200 -- the render_rays hot path of Freedomcls/nerf-siren on B200 (sm_100a).

Host side: PyTorch modules / functions with the reference's names and signatures
(``models.rendering.render_rays``, ``models.nerf.{Embedding,NeRF,SemanticNeRF}``, ``models.nerf_cls.NeRF_3D``).
Device side: hand-written CUDA kernels behind a C ABI (``include/nerf_b200.h`` -> ``libnerf_b200.so``).
Importing this package loads the shared library and fails loudly if it has not been built; nothing in the
package falls back to torch ops or to the CPU.
"""
import sys as _sys

from . import _lib  # noqa: F401  (raises ImportError when libnerf_b200.so is missing)
from . import config, dp, eg3d, evaluation, losses, ops, synth, system  # noqa: F401
from .config import get_precision, set_precision  # noqa: F401
from .models import nerf, nerf_cls, rendering  # noqa: F401
from .models.nerf import Embedding, NeRF, SemanticNeRF, SirenField  # noqa: F401
from .models.nerf_cls import NeRF_3D  # noqa: F401
from .models.rendering import render_rays, render_rays_3d, render_rays_3d_conv, sample_pdf  # noqa: F401

__version__ = "0.1.0"


def install_as_models() -> None:
    """Make ``from models.rendering import render_rays`` / ``from models.nerf import *`` (system.py:5-8,
    eval.py:9-10 in the reference) resolve to this package: the drop-in switch for the reference's callers."""
    from . import models as _m
    _sys.modules["models"] = _m
    _sys.modules["models.rendering"] = rendering
    _sys.modules["models.nerf"] = nerf
    _sys.modules["models.nerf_cls"] = nerf_cls


def install_as_losses() -> None:
    """Make ``from losses import loss_dict`` (system.py:10, train.py in the reference) resolve to this package."""
    _sys.modules["losses"] = losses


def install_as_eg3d() -> None:
    """Make ``from volumetric_rendering.renderer import ImportanceRenderer`` / ``...ray_marcher import MipRayMarcher2``
    (eg3d_training/triplane.py:15-16 in the reference) resolve to this package's kernel-backed mirror."""
    import types
    pkg = types.ModuleType("volumetric_rendering")
    pkg.renderer = eg3d
    pkg.ray_marcher = eg3d
    _sys.modules["volumetric_rendering"] = pkg
    _sys.modules["volumetric_rendering.renderer"] = eg3d
    _sys.modules["volumetric_rendering.ray_marcher"] = eg3d
