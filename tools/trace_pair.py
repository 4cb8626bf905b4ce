"""Diagnostics: clock64 timeline of block 0 of the tile-pair forward (first 64 events = (step, slot) pairs).  Needs a library
built with NERFB_NVCC_EXTRA=-DNERFB_TRACE_BUILD.  Per event: MMA thread [start, +wait epi_done, slab-wait cycles, +end],
epilogue warp 0 [start, +wait acc_full, +drained]."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
save = len(sys.argv) > 1 and sys.argv[1] == "train"
os.environ["NERFB_FWD_PAIR"] = "3" if save else "1"
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
buf = torch.zeros(64 * 16, dtype=torch.int64, device=dev)
ws = ops.mlp_workspace(desc, _lib.PREC_BF16, n * S, False, save, dev)
with torch.no_grad():
    for _ in range(3):
        ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, None, rays, z, n, S, False, save, ws)
    torch.cuda.synchronize()
    _lib.lib.nerfb_debug_trace.argtypes = [C.c_void_p]
    _lib.lib.nerfb_debug_trace(buf.data_ptr())
    ops.mlp_fwd(desc, _lib.PREC_BF16, _lib.IN_RAYS, flat, None, rays, z, n, S, False, save, ws)
    torch.cuda.synchronize()
    _lib.lib.nerfb_debug_trace(None)
t = buf.cpu().view(64, 16)
t0 = int(t[0, 0])
for e in range(64):
    r = [int(v) for v in t[e]]
    print(f"ev {e:2d} step {(e // 2) % 10} slot {e % 2} | mma start {r[0]-t0:7d} wait_epi +{r[1]-r[0] if r[1] else 0:5d} slab_wait {r[2]:5d} dur {r[3]-r[0]:5d} | "
          f"epi start {r[8]-t0:7d} wait_acc +{r[9]-r[8]:5d} drain +{r[10]-r[9]:5d} (of which store-counter wait {r[11]:5d})")
