"""Convergence check of the whole training path on a scene with a closed-form image: a shaded unit sphere on a white
background (Blender-shaped cameras, near 2 / far 6).  Trains the coarse + fine NeRF pair with the captured data-parallel
step for --steps iterations of 4096 random rays and reports training loss and held-out-view PSNR (through
evaluation.batched_inference) as JSON lines.     python tools/train_demo.py [--steps 2000] [--model nerf|siren]"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import evaluation as ev, synth  # noqa: E402


def sphere_rgb(rays):
    """Lambert-shaded unit sphere at the origin, white background: the pixel colour each ray should produce."""
    o, d = rays[:, :3], rays[:, 3:6]
    b = (o * d).sum(-1)
    disc = b * b - ((o * o).sum(-1) - 1.0)
    hit = disc > 0
    t = -b - torch.sqrt(disc.clamp_min(0))
    n = torch.nn.functional.normalize(o + d * t[:, None], dim=-1)
    light = torch.nn.functional.normalize(torch.tensor([0.4, 0.3, 0.85], device=rays.device), dim=0)
    shade = (n @ light).clamp_min(0) * 0.8 + 0.2
    col = (0.5 + 0.5 * n) * shade[:, None]
    return torch.where(hit[:, None], col, torch.ones_like(col))


ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=2000)
ap.add_argument("--rays", type=int, default=4096)
ap.add_argument("--model", default="nerf", choices=["nerf", "siren"])
ap.add_argument("--precision", default="auto", choices=["auto", "fp32"])
ap.add_argument("--lr", type=float, default=5e-4)
args = ap.parse_args()
dev = torch.device("cuda:0")
torch.manual_seed(0)
nb.set_precision(args.precision)
if args.model == "siren":
    models = [nb.SirenField().to(dev), nb.SirenField().to(dev)]
    emb = [nb.Embedding(3, 0), nb.Embedding(3, 0)]
else:
    models = [nb.NeRF().to(dev), nb.NeRF().to(dev)]
    emb = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
trainer = nb.dp.DataParallelStep(models, emb, 64, 64, False, 1.0, 1.0, 32768, True, lr=args.lr, cuda_graph=True)
pool = [synth.blender_rays(args.rays, seed=100 + i).to(dev) for i in range(64)]
targets = [sphere_rgb(r) for r in pool]
held = synth.blender_image_rays(200, 200, seed=987654).to(dev)          # a view no training batch draws from
held_rgb = sphere_rgb(held)


def psnr_held():
    out = ev.batched_inference(models, emb, held, 64, 64, False, 32768, True)
    return float(-10 * torch.log10(((out["rgb_fine"] - held_rgb) ** 2).mean()))


print(json.dumps({"step": 0, "held_out_psnr": psnr_held()}), flush=True)
t0 = time.time()
for step in range(1, args.steps + 1):
    loss = trainer.step(pool[step % len(pool)], targets[step % len(pool)])
    if step % max(1, args.steps // 10) == 0:
        torch.cuda.synchronize()
        print(json.dumps({"step": step, "loss": float(loss), "held_out_psnr": psnr_held(), "wall_s": round(time.time() - t0, 2)}), flush=True)
