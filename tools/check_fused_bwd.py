"""Diagnostic: fused backward (chain + wgrad CTAs in ONE launch, dZ through an L2-resident ring) against the 3-launch
backward (dgrad -> HBM record -> wgrad) on the bench shapes: per-parameter relative difference and CUDA-event timings.
Usage: python tools/check_fused_bwd.py [nerf|siren|cls] [n_rays] [S] ["G,Q;G,Q;..."]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import _field, _lib, ops, synth  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "nerf"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
S = int(sys.argv[3]) if len(sys.argv) > 3 else 128
combos = [tuple(int(v) for v in c.split(",")) for c in (sys.argv[4] if len(sys.argv) > 4 else "8,10").split(";")]   # G,Q[,budget,C1,C0,CN]
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = {"nerf": nb.NeRF, "siren": nb.SirenField, "cls": lambda: nb.NeRF_3D(CLS=6)}[kind]().to(dev)
siren = kind == "siren"
desc = m._desc(False, 0, 0, True) if siren else m._desc(False, 10, 4, True)
film = m._film() if siren else None
rays = synth.blender_rays(n, seed=1).to(dev)
z = torch.sort(torch.rand(n, S, device=dev) * 4 + 2, -1)[0].contiguous()
flat = _field.flat_params(m).detach()


def run(fused, G=8, Q=10, *extra, iters=10):
    os.environ["NERFB_BWD_FUSED"] = "1" if fused else "0"
    os.environ["NERFB_FUSED_G"], os.environ["NERFB_FUSED_Q"] = str(G), str(Q)
    for name, val in zip(("NERFB_FUSED_BUDGET", "NERFB_FUSED_C1", "NERFB_FUSED_C0", "NERFB_FUSED_CN"), extra):
        os.environ[name] = str(val)
    with torch.no_grad():
        ws = ops.mlp_workspace(desc, _lib.PREC_BF16, n * S, False, True, dev)
        raw = ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n, S, False, True, ws)
        d_raw = torch.randn(raw.shape, device=dev, generator=torch.Generator(device=dev).manual_seed(5)) * 1e-3
        g, _ = ops.mlp_bwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n, S, raw, d_raw, ws)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            ops.mlp_bwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, film, rays, z, n, S, raw, d_raw, ws)
        e1.record()
        torch.cuda.synchronize()
    return g, e0.elapsed_time(e1) / iters, ws.numel()


g0, t0, w0 = run(False)
print(json.dumps({"kind": kind, "rays": n, "S": S, "mode": "3-launch", "ms": round(t0, 4), "workspace_GB": round(w0 / 1e9, 3)}), flush=True)
for combo in combos:
    G, Q = combo[0], combo[1]
    g1, t1, w1 = run(True, *combo)
    worst, off = (None, 0.0), 0
    for k, p in m.named_parameters():
        a, b = g0[off:off + p.numel()].double(), g1[off:off + p.numel()].double()
        off += p.numel()
        rel = ((a - b).norm() / (a.norm() + 1e-30)).item()
        if rel > worst[1]:
            worst = (k, rel)
    print(json.dumps({"kind": kind, "mode": "fused", "G": G, "Q": Q, "extra": list(combo[2:]), "ms": round(t1, 4), "workspace_GB": round(w1 / 1e9, 3),
                      "worst_rel_diff_vs_3launch": worst, "finite": bool(torch.isfinite(g1).all())}), flush=True)
