"""Profiling driver: runs the MLP kernels alone on the bench shapes (4096 rays x 128 fine samples) so that
`ncu --set full -k regex:mlp_tc` captures them without the rest of the step.  Usage: python tools/prof_mlp.py [fwd|train|siren|siren_train] [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import _field, _lib, ops, synth  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "fwd"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
torch.manual_seed(0)
siren = mode.startswith("siren")
m = (nb.SirenField() if siren else nb.NeRF()).to(dev)
desc = m._desc(False, 0, 0, True) if siren else m._desc(False, 10, 4, True)
film = m._film() if siren else None
n, S = 4096, 128
rays = synth.blender_rays(n, seed=1).to(dev)
z = torch.sort(torch.rand(n, S, device=dev) * 4 + 2, -1)[0].contiguous()
flat = _field.flat_params(m).detach()
save = mode in ("train", "siren_train")
with torch.no_grad():
    ws = ops.mlp_workspace(desc, _lib.PREC_BF16, n * S, False, save, dev)
    for _ in range(iters):
        raw = ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n, S, False, save, ws)
        if save:
            d_raw = torch.randn_like(raw) * 1e-3
            ops.mlp_bwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n, S, raw, d_raw, ws)
    torch.cuda.synchronize()
print("ok", float(raw.abs().mean()))
