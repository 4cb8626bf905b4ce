"""Diagnostics: clock64 timeline of block 0 of the tcgen05 forward kernel (first 40 steps = 4 tiles).
Per step: MMA thread [t_kb_start x5 | t_before_w_full x5], epilogue [group0: acc_full seen, block done x2 | group1: same]."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import _field, _lib, ops, synth  # noqa: E402

save = len(sys.argv) > 1 and sys.argv[1] == "train"
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = nb.NeRF().to(dev)
desc = m._desc(False, 10, 4, True)
n, S = 4096, 128
rays = synth.blender_rays(n, seed=1).to(dev)
z = torch.sort(torch.rand(n, S, device=dev) * 4 + 2, -1)[0].contiguous()
flat = _field.flat_params(m).detach()
buf = torch.zeros(40 * 32, dtype=torch.int64, device=dev)
ws = ops.mlp_workspace(desc, _lib.PREC_BF16, n * S, False, save, dev)
with torch.no_grad():
    ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, None, rays, z, n, S, False, save, ws)
    torch.cuda.synchronize()
    _lib.lib.nerfb_debug_trace.argtypes = [C.c_void_p]
    _lib.lib.nerfb_debug_trace(buf.data_ptr())
    ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, None, rays, z, n, S, False, save, ws)
    torch.cuda.synchronize()
    _lib.lib.nerfb_debug_trace(None)
t = buf.cpu().view(40, 32)
t0 = int(t[0, 0])
for g in range(10, 24):
    r = [int(v) - t0 if int(v) else -1 for v in t[g]]
    kb = " ".join(f"[{r[4*k]} a+{r[4*k+1]-r[4*k]} w+{r[4*k+2]-r[4*k+1]} i+{r[4*k+3]-r[4*k+2]}]" for k in range(5) if r[4 * k] >= 0)
    e = r[20]
    print(f"gs {g:2d} | {kb} | epi acc {e} +ld {r[21]-e} +ld2 {r[26]-e} +math {r[27]-e} +sts {r[22]-e} +fence {r[23]-e} +arr0 {r[24]-e} +arr1 {r[25]-e}")
