"""EG3D training step at BASELINE config 5's shape (3 x 32 x 256^2 planes, 16,384 rays, 48 + 48 samples): CUDA-event times of
the differentiable schedule of nerf_siren_b200.eg3d (forward, backward, the two backward kernels alone) beside the fused
no_grad render.  Usage: python tools/bench_eg3d_train.py [rays] [res]"""
import sys

import torch

sys.path.insert(0, ".")
from nerf_siren_b200 import eg3d, ops            # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
RES = int(sys.argv[2]) if len(sys.argv) > 2 else 256
DEV = "cuda"


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


torch.manual_seed(0)
kw = dict(eg3d.DEFAULT_RENDERING_KWARGS, depth_resolution=48, depth_resolution_importance=48, ray_start=2.2, ray_end=3.3, box_warp=1.0)
ren = eg3d.EG3D_Renderer(eg3d.LearnedPlanes(resolution=RES), kw).to(DEV)
g = torch.Generator(device=DEV).manual_seed(1)
d = torch.nn.functional.normalize(torch.randn(R, 3, device=DEV, generator=g) * 0.15 + torch.tensor([0, 0, -1.0], device=DEV), dim=-1)
o = torch.tensor([0, 0, 2.7], device=DEV).expand(R, 3).contiguous() + torch.randn(R, 3, device=DEV, generator=g) * 0.01
target = torch.rand(R, 3, device=DEV, generator=g)


def render_nograd():
    with torch.no_grad():
        ren.render(-1, o, d)


def fwd():
    out = ren.render(-1, o, d)
    return ((out["rgb_coarse"] - target) ** 2).mean() + ((out["rgb_fine"] - target) ** 2).mean()


def step():
    for p in ren.parameters():
        p.grad = None
    fwd().backward()


P = R * 96
cl = ops.triplane_relayout(ren.plane_source.planes.detach().view(3, 32, RES, RES))
dec = ren.decoder.flat()
coords = (torch.rand(P, 3, device=DEV, generator=g) - 0.5) * 0.9
d_rgb, d_sigma = torch.randn(P, 3, device=DEV, generator=g), torch.randn(P, device=DEV, generator=g)
colors, dens = torch.rand(R, 96, 3, device=DEV, generator=g), torch.randn(R, 96, device=DEV, generator=g)
depths = torch.sort(torch.rand(R, 96, device=DEV, generator=g) + 2.2, dim=1)[0].contiguous()
g_rgb = torch.randn(R, 3, device=DEV, generator=g)

with torch.no_grad():
    smp = ops.triplane_render(cl, dec, 1.0, o, d, 48, 48, 2.2, 3.3, None, None, False, 1.0, False, None, None, 1, return_samples=True)[6]
ray_coords = torch.addcmul(o.unsqueeze(1), smp[:, 0].unsqueeze(-1), d.unsqueeze(1)).reshape(-1, 3).contiguous()
t_inf = timed(render_nograd)
t_fwd = timed(lambda: fwd())
t_step = timed(step)
t_field = timed(lambda: ops.triplane_field(cl, dec, 1.0, coords, 1.0))
t_fbwd = timed(lambda: ops.triplane_field_bwd(cl, dec, 1.0, coords, 1.0, d_rgb, d_sigma))
t_fbwd_ray = timed(lambda: ops.triplane_field_bwd(cl, dec, 1.0, ray_coords, 1.0, d_rgb, d_sigma))
t_mbwd = timed(lambda: ops.mip_march_bwd(colors, dens, depths, False, g_rgb, None, None))
print(f"rays {R} planes {RES}^2 samples 48+48 ({P} points)")
print(f"  fused no_grad render            {t_inf:8.3f} ms  ({R / t_inf / 1e3:.2f} M rays/s)")
print(f"  differentiable forward          {t_fwd:8.3f} ms")
print(f"  forward + backward              {t_step:8.3f} ms  ({R / t_step / 1e3:.2f} M rays/s)")
print(f"  nerfb_triplane_field  ({P} pts) {t_field:8.3f} ms")
print(f"  nerfb_triplane_field_bwd        {t_fbwd:8.3f} ms  on uniformly random points ({P * 12 * 128 / t_fbwd / 1e6:.1f} GB/s of 128 B plane-gradient reductions)")
print(f"  nerfb_triplane_field_bwd        {t_fbwd_ray:8.3f} ms  on the render's own sample points (ray-coherent, unsorted: S coarse then Si importance)")
print(f"  nerfb_mip_march_bwd ({R} x 96)  {t_mbwd:8.3f} ms")
