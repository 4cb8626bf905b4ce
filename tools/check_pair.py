"""Diagnostic: tile-pair forward (mlp_tc_fwd_pair_kernel, NERFB_FWD_PAIR=1) against the single-tile forward on the bench
shapes: max-abs output difference and CUDA-event timings.  Usage: python tools/check_pair.py [n_rays] [iters]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import _field, _lib, ops, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
modes = [int(v) for v in (sys.argv[3] if len(sys.argv) > 3 else "1,2").split(",")]
dev = torch.device("cuda:0")


def run(kind, S, sigma_only, pair, n_rays):
    os.environ["NERFB_FWD_PAIR"] = str(int(pair))
    torch.manual_seed(0)
    m = {"nerf": nb.NeRF, "siren": nb.SirenField, "cls": lambda: nb.NeRF_3D(CLS=6)}[kind]().to(dev)
    siren = kind == "siren"
    desc = m._desc(sigma_only, 0, 0, True) if siren else m._desc(sigma_only, 10, 4, True)
    film = m._film() if siren else None
    rays = synth.blender_rays(n_rays, seed=1).to(dev)
    if siren:
        rays[:, :6] *= 0.05
    z = torch.sort(torch.rand(n_rays, S, device=dev, generator=torch.Generator(device=dev).manual_seed(3)) * 4 + 2, -1)[0].contiguous()
    flat = _field.flat_params(m).detach()
    with torch.no_grad():
        ws = ops.mlp_workspace(desc, _lib.PREC_BF16, n_rays * S, sigma_only, False, dev)
        out = ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n_rays, S, sigma_only, False, ws)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n_rays, S, sigma_only, False, ws)
        e1.record()
        torch.cuda.synchronize()
    return out.clone(), e0.elapsed_time(e1) / iters


for kind, S, so, nr in (("nerf", 128, False, n), ("nerf", 64, False, n), ("nerf", 64, True, n), ("nerf", 128, False, 1037), ("nerf", 64, False, 5),
                        ("siren", 128, False, n), ("siren", 64, True, n), ("cls", 128, False, n)):
    try:
        a, ta = run(kind, S, so, 0, nr)
        for mode in modes:
            b, tb = run(kind, S, so, mode, nr)
            d = (a - b).abs()
            print(json.dumps({"kind": kind, "rays": nr, "S": S, "sigma_only": so, "mode": mode, "single_ms": round(ta, 4), "pair_ms": round(tb, 4),
                              "speedup": round(ta / tb, 3), "max_abs_diff": float(d.max()), "mean_abs_diff": float(d.mean()),
                              "out_absmax": float(a.abs().max()), "finite": bool(torch.isfinite(b).all())}), flush=True)
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"kind": kind, "S": S, "sigma_only": so, "error": str(e)[:300]}), flush=True)
