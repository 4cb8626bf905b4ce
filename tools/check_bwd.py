"""Diagnostic: per-parameter relative error of the bf16 tensor-core backward against the fp32 CUDA-core arm."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_siren_b200 as nb
from nerf_siren_b200 import _field, _lib, ops, synth
dev = torch.device("cuda:0")
torch.manual_seed(0)
n, S = int(sys.argv[1]) if len(sys.argv) > 1 else 300, 64
gain = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
m = nb.NeRF().to(dev)
with torch.no_grad():
    for k, p in m.named_parameters():
        if k.endswith("weight"): p.mul_(gain)
shift = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0        # > 0: every ReLU gate open (network is linear: no gate flips)
with torch.no_grad():
    for k, p in m.named_parameters():
        if k.endswith("bias") and ("xyz_encoding_" in k or "dir_encoding" in k) and "final" not in k: p.add_(shift)
desc = m._desc(False, 10, 4, True)
rays = synth.blender_rays(n, seed=1).to(dev)
z = torch.sort(torch.rand(n, S, device=dev) * 4 + 2, -1)[0].contiguous()
flat = _field.flat_params(m).detach()
d_raw = torch.randn(n * S, 4, device=dev)
res = {}
for prec in (_lib.PREC_FP32, _lib.PREC_BF16):
    ws = ops.mlp_workspace(desc, prec, n * S, False, True, dev)
    raw = ops.mlp_fwd(desc, prec, _lib.IN_RAYS, flat, None, rays, z, n, S, False, True, ws)
    g, _ = ops.mlp_bwd(desc, prec, _lib.IN_RAYS, flat, None, rays, z, n, S, raw, d_raw, ws)
    torch.cuda.synchronize()
    res[prec] = (raw, g)
print("fwd max err", (res[0][0] - res[1][0]).abs().max().item(), "scale", res[0][0].abs().max().item())
off = 0
for k, p in m.named_parameters():
    a, b = res[0][1][off:off + p.numel()], res[1][1][off:off + p.numel()]
    off += p.numel()
    print(f"{k:28s} ref_norm {a.norm().item():.4e} rel_err {((a - b).norm() / (a.norm() + 1e-30)).item():.4e}")
# against the bf16-emulating oracle (CPU)
from oracle import render_oracle as O
rc, zc = rays.cpu(), z.cpu()
xyz = (rc[:, None, :3] + rc[:, None, 3:6] * zc[..., None]).reshape(-1, 3)
x = torch.cat([O.embed(xyz, 10), torch.repeat_interleave(O.embed(rc[:, 3:6], 4), S, 0)], -1)
p = {k: v.detach().cpu() for k, v in m.state_dict().items()}
out_e, g_e = O.nerf_mlp_bf16(p, x, d_raw.cpu())
print("vs bf16 emulation: fwd max err", (res[1][0].cpu() - out_e).abs().max().item())
off = 0
for k, pp in m.named_parameters():
    b = res[1][1][off:off + pp.numel()].cpu().view_as(pp); off += pp.numel()
    print(f"{k:28s} rel_err_vs_emulation {((g_e[k] - b).norm() / (g_e[k].norm() + 1e-30)).item():.4e}")
