"""BASELINE config 5: EG3D tri-plane renderer, 3 x 32 x 256^2 planes, 128^2 = 16,384 rays, 48 + 48 samples per ray.
Times kernel (6) (CUDA events on the launch stream, planes already channels-last and resident) and, with --cpu, the
oracle port of the reference's ImportanceRenderer on the host cores.  Prints one JSON line.
    python tools/bench_eg3d.py [--rays 16384] [--iters 50] [--cpu]"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

ap = argparse.ArgumentParser()
ap.add_argument("--rays", type=int, default=16384)
ap.add_argument("--iters", type=int, default=50)
ap.add_argument("--cpu", action="store_true")
args = ap.parse_args()

from oracle import eg3d_oracle as E  # noqa: E402  (input generators + the CPU baseline leg only)
import nerf_siren_b200 as nb  # noqa: E402

dev = torch.device("cuda:0")
S = Si = 48
planes = E.planes_init(seed=5, res=256)
decp = E.decoder_init(seed=6)
dec = torch.cat([decp["net.0.weight"].reshape(-1), decp["net.0.bias"], decp["net.2.weight"].reshape(-1), decp["net.2.bias"]]).to(dev)
ro, rd = E.camera_rays(args.rays, seed=7)
cl = nb.ops.triplane_relayout(planes[0].to(dev))
o, d = ro[0].to(dev), rd[0].to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run(seed):
    return nb.ops.triplane_render(cl, dec, 1.0, o, d, S, Si, 2.2, 3.3, None, None, False, 1.0, False, None, None, seed)


for i in range(5):
    run(i)
torch.cuda.synchronize()
ts = []
for i in range(args.iters):
    flush.zero_()                                      # evict the planes from L2 between iterations (cold planes, warm code)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(100 + i); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ts.sort()
ms = ts[len(ts) // 2]
samples = args.rays * (S + Si)
line = {"metric": "rays/s, EG3D tri-plane render (48+48 samples, 3x32x256^2 planes)", "value": args.rays / ms * 1e3, "unit": "rays/s",
        "ms": ms, "rays": args.rays, "gather_gbs": samples * 1536 / ms / 1e6, "gather_bytes_per_sample": 1536,
        "decoder_gflops": samples * 2 * 2304 / ms / 1e6, "launches_per_call": 3,
        "l2": "256 MB flush between iterations"}
if args.cpu:
    n = 1024
    opts = {"depth_resolution": S, "depth_resolution_importance": Si, "ray_start": 2.2, "ray_end": 3.3, "box_warp": 1.0,
            "white_back": False, "clamp_mode": "softplus"}
    torch.set_num_threads(os.cpu_count())
    g = torch.Generator().manual_seed(1)
    rc, u = torch.rand(1, n, S, 1, generator=g), torch.rand(n, Si, generator=g)
    with torch.no_grad():
        E.importance_render(planes, decp, ro[:, :n], rd[:, :n], opts, rc, u)
        t0 = time.perf_counter()
        for _ in range(3):
            E.importance_render(planes, decp, ro[:, :n], rd[:, :n], opts, rc, u)
        dt = (time.perf_counter() - t0) / 3
    line["cpu_baseline"] = {"value": n / dt, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "port",
                            "sample": f"{n} rays x (48 + 48) samples, 3 timed calls after 1 warm-up ({dt * 1e3:.0f} ms/call)"}
print(json.dumps(line))
