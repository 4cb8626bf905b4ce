"""Kernel (1) forward alone, inference vs save-for-backward, CUDA-event timed (4096 rays x 128 samples).
Usage: python tools/time_fwd.py [nerf|siren] [tag]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import _field, _lib, ops, synth  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "nerf"
tag = sys.argv[2] if len(sys.argv) > 2 else ""
dev = torch.device("cuda:0")
torch.manual_seed(0)
siren = kind == "siren"
m = (nb.SirenField() if siren else nb.NeRF()).to(dev)
desc = m._desc(False, 0, 0, True) if siren else m._desc(False, 10, 4, True)
film = m._film() if siren else None
n, S = 4096, 128
rays = synth.blender_rays(n, seed=1).to(dev)
z = torch.sort(torch.rand(n, S, device=dev) * 4 + 2, -1)[0].contiguous()
flat = _field.flat_params(m).detach()
out = {}
with torch.no_grad():
    for save in (False, True):
        ws = ops.mlp_workspace(desc, _lib.PREC_BF16, n * S, False, save, dev)
        for _ in range(5):
            ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n, S, False, save, ws)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30):
            ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n, S, False, save, ws)
        e1.record()
        torch.cuda.synchronize()
        out["save" if save else "infer"] = round(e0.elapsed_time(e1) / 30, 4)
print(kind, tag, out, flush=True)
