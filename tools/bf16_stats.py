"""Diagnostic: error statistics of the bf16 arm against the reference fixtures (end-to-end render_rays)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb
from tests._util import CASES, gold, rng_of
from tests.test_gpu_render import _ReplayRNG, _models_for
nb.set_precision("bf16")
for name in sorted(CASES):
    if name == "llff_cls": continue
    g = gold("g9_" + name); cfg = CASES[name]
    mc, mf = _models_for(name)
    emb = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
    with _ReplayRNG(rng_of(g), cfg), torch.no_grad():
        out = nb.render_rays([mc, mf], emb, g["rays"].cuda(), cfg["n_samples"], cfg["use_disp"], cfg["perturb"], cfg["noise_std"],
                 cfg["n_importance"], cfg["chunk"], cfg["white_back"], test_time=cfg["test_time"])
    for k in sorted(out):
        ref = g["out:" + k]; e = (out[k].cpu() - ref).abs().flatten()
        mse = ((out[k].cpu() - ref) ** 2).mean().item()
        print(f"{name:20s} {k:15s} max {e.max():.4f} mean {e.mean():.5f} p99 {e.kthvalue(int(0.99*e.numel()))[0]:.4f} psnr {-10*torch.log10(torch.tensor(mse+1e-20)).item():.1f}")
