"""Drop-in for the reference's ``models/nerf_cls.py`` NeRF_3D (:12-112): NeRF plus the per-sample class head
``parse = Linear(W, 2W) -> Linear(2W, CLS) -> Sigmoid`` on xyz_encoding_final; output (B, 4 + CLS).
The reference's NeRF_Point (:116-203) is dead code (bad super() call) and is not provided; importing this
module needs neither MinkowskiEngine nor open3d."""
from torch import nn

from .nerf import Embedding, NeRF  # noqa: F401  (the reference re-exports Embedding from here)

__all__ = ["NeRF_3D"]


class NeRF_3D(NeRF):
    def __init__(self, D=8, W=256, CLS=3, in_channels_xyz=63, in_channels_dir=27, skips=[4]):
        super().__init__(D=D, W=W, in_channels_xyz=in_channels_xyz, in_channels_dir=in_channels_dir, skips=skips)
        self.CLS = CLS
        self.parse = nn.Sequential(nn.Linear(W, 2 * W), nn.Linear(2 * W, CLS), nn.Sigmoid())
