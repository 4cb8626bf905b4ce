"""Diagnostics: clock64 timeline of ONE wgrad CTA (unit 0, index 0) of the fused backward, first 120 tiles.  Needs a library built
with NERFB_NVCC_EXTRA=-DNERFB_TRACE_BUILD.  Producer lane per tile: [start, +flag seen, +fence, +stage free, +copies issued];
MMA lane: [wait for the first half, +landed, ..., both halves issued]."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["NERFB_BWD_FUSED"] = "1"
os.environ["NERFB_FUSED_G"] = sys.argv[1] if len(sys.argv) > 1 else "6"
os.environ["NERFB_FUSED_Q"] = sys.argv[2] if len(sys.argv) > 2 else "12"
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import _field, _lib, ops, synth  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nb.NeRF().to(dev)
desc = m._desc(False, 10, 4, True)
n, S = 4096, 128
rays = synth.blender_rays(n, seed=1).to(dev)
z = torch.sort(torch.rand(n, S, device=dev) * 4 + 2, -1)[0].contiguous()
flat = _field.flat_params(m).detach()
buf = torch.zeros(120 * 8 + 150 * 8 + 148 * 8, dtype=torch.int64, device=dev)
with torch.no_grad():
    ws = ops.mlp_workspace(desc, _lib.PREC_BF16, n * S, False, True, dev)
    raw = ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, None, rays, z, n, S, False, True, ws)
    d_raw = torch.randn_like(raw) * 1e-3
    for _ in range(2):
        ops.mlp_bwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, None, rays, z, n, S, raw, d_raw, ws)
    torch.cuda.synchronize()
    _lib.lib.nerfb_debug_trace_bwd.argtypes = [C.c_void_p]
    _lib.lib.nerfb_debug_trace_bwd(buf.data_ptr())
    ops.mlp_bwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, None, rays, z, n, S, raw, d_raw, ws)
    torch.cuda.synchronize()
    _lib.lib.nerfb_debug_trace_bwd(None)
full = buf.cpu()
t = full[:120 * 8].view(120, 8)
t0 = int(t[0, 0])
prev = t0
show = int(sys.argv[3]) if len(sys.argv) > 3 else 40
for e in range(min(120, show)):
    r = [int(v) for v in t[e]]
    if r[0] == 0:
        break
    print(f"tile {e:3d} | prod start {r[0]-t0:8d} (+{r[0]-prev:6d}) flag +{r[1]-r[0]:6d} fence +{r[2]-r[1]:5d} stage +{r[3]-r[2]:6d} issued +{r[4]-r[3]:5d} | "
          f"mma wait@ {r[5]-t0:8d} landed +{r[6]-r[5]:6d} done +{r[7]-r[6]:6d}")
    prev = r[0]

c = full[120 * 8:120 * 8 + 150 * 8].view(150, 8)
tot = full[120 * 8 + 150 * 8:].view(148, 8)
print('wgrad CTAs: unit items tiles | total clk | flag-wait clk | stage-wait clk | busy per tile')
for wi in range(148):
    r = [int(v) for v in tot[wi]]
    if r[0] == 0:
        continue
    print(f'w {wi:3d} unit {r[3]:2d} items {r[4]} tiles {r[5]:4d} | {r[0]:9d} | {r[1]:9d} | {r[2]:9d} | {(r[0]-r[1]-r[2])/max(r[5],1):8.0f}')
c0 = int(c[0, 0])
prev = c0
print("chain CTA 0, store lane, per packet: start | slot-free wait | wait for blocks (sum) | publish waits (sum) | total")
for e in range(min(150, 3 * show)):
    r = [int(v) for v in c[e]]
    if r[0] == 0:
        break
    print(f"pkt {e:3d} (tile {e // 11} step {e % 11}) | start {r[0]-c0:8d} (+{r[0]-prev:6d}) slot +{r[1]-r[0]:6d} blocks {r[2]:6d} publish {r[3]:6d} total {r[4]-r[0]:6d}")
    prev = r[0]
