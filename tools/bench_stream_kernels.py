"""Kernels (2)(3)(4) at full-image size (800 x 800 = 640,000 rays, 64 coarse / 128 fine samples): achieved HBM GB/s of the
algorithmic bytes against the measured copy peak.  At the 4096-ray training batch these kernels move 2-10 MB
and are launch-latency-bound; this is the bandwidth-bound regime (full-image render, eval.py).  One JSON line per kernel.
    python tools/bench_stream_kernels.py [--rays 640000]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import ops, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rays", type=int, default=640000)
args = ap.parse_args()
dev = torch.device("cuda:0")
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
n, S, Si = args.rays, 64, 64
Sf = S + Si
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters=10):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


rays = synth.blender_image_rays(800, 800, seed=1)[:n].contiguous().to(dev)
n = rays.shape[0]
pr = torch.rand(n, S, device=dev)
z = ops.sample_coarse(rays, S, False, 1.0, pr)
raw = torch.rand(n, S, 4, device=dev); raw[..., 3] = raw[..., 3] * 6 - 2
noise = torch.randn(n, S, device=dev)
w = ops.composite_fwd(raw, z, rays, noise, 1.0, True)[0]
u = torch.rand(n, Si, device=dev)
zf = ops.resample(z, w, Si, False, u)
rawf = torch.rand(n, Sf, 4, device=dev)
noisef = torch.randn(n, Sf, device=dev)
d_rgb = torch.randn(n, 3, device=dev)
rawf_copy = torch.empty_like(rawf)
cases = [
    ("sample_coarse_kernel", lambda: ops.sample_coarse(rays, S, False, 1.0, pr), n * (32 + 8 * S)),
    ("composite_fwd_kernel (64 samples)", lambda: ops.composite_fwd(raw, z, rays, noise, 1.0, True), n * S * 28 + n * 32),
    ("composite_fwd_kernel (128 samples)", lambda: ops.composite_fwd(rawf, zf, rays, noisef, 1.0, True), n * Sf * 28 + n * 32),
    ("composite_bwd_kernel (128 samples)", lambda: ops.composite_bwd(rawf, zf, rays, noisef, 1.0, True, d_rgb=d_rgb), n * Sf * 40 + n * 24),
    ("resample_kernel", lambda: ops.resample(z, w, Si, False, u), n * (8 * S + 4 * Si + 4 * Sf)),
    ("resample_kernel (det: render path)", lambda: ops.resample(z, w, Si, True, None), n * (8 * S + 4 * Sf)),
    ("generate_rays_kernel", lambda: ops.generate_rays(torch.eye(3, 4), 800, 800, 1111.1), 640000 * 32),
    # what the memory system gives torch's own streaming kernels at the SAME sizes (L2 flushed): the ceiling a write-only
    # kernel like sample_coarse (164 MB out, 20 MB in) or a 2:1 read:write kernel can be held against
    ("reference: torch fill_ (write-only, size of the z output)", lambda: z.fill_(1.0), z.numel() * 4),
    ("reference: torch copy_ (read + write, size of raw 128 samples)", lambda: rawf_copy.copy_(rawf), 2 * rawf.numel() * 4),
]
for name, fn, by in cases:
    ms = timed(fn)
    print(json.dumps({"kernel": name, "rays": n, "ms": ms, "bytes": by, "achieved": by / ms / 1e6, "peak": peak, "unit": "GB/s",
                      "frac": by / ms / 1e6 / peak, "bound": "hbm", "l2": "512 MB flush before every timed launch"}), flush=True)
