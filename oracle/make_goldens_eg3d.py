"""Pin oracle/eg3d_oracle.py against the REAL reference (volumetric_rendering/, eg3d_training/triplane.py) and emit
tests/golden/g10_eg3d_*.npz.  Build container only (needs /root/reference; imported, never copied).  Stored outputs
are the REFERENCE's.  TEST INFRASTRUCTURE.

    python oracle/make_goldens_eg3d.py
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
REF = os.environ.get("NERF_REFERENCE", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, REF)

from oracle import eg3d_oracle as E  # noqa: E402
from oracle.make_goldens import replay_rng  # noqa: E402


def same(a, b, what):
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert torch.equal(a, b), (what, (a - b).abs().max().item())


def main():
    from volumetric_rendering.renderer import ImportanceRenderer, sample_from_planes, generate_planes
    from volumetric_rendering.ray_marcher import MipRayMarcher2
    from eg3d_training.triplane import OSGDecoder

    torch.manual_seed(0)
    cases = {
        # name: (plane res, rays, S, Si, ray_start, ray_end, box_warp, white_back, disparity)
        "small": (64, 256, 16, 16, 2.25, 3.3, 1.0, False, False),
        "white": (64, 192, 24, 24, 2.0, 3.4, 1.2, True, False),
        "disp": (32, 128, 12, 20, 2.0, 3.5, 1.5, False, True),
        "cfg5": (256, 512, 48, 48, 2.2, 3.3, 1.0, False, False),
    }
    for name, (res, R, S, Si, rs, re, bw, wb, disp) in cases.items():
        planes = E.planes_init(seed=11, res=res)
        decp = E.decoder_init(seed=12)
        dec = OSGDecoder(32, {"decoder_lr_mul": 1, "decoder_output_dim": 3})
        dec.load_state_dict(decp)
        ro, rd = E.camera_rays(R, seed=13)
        opts = {"depth_resolution": S, "depth_resolution_importance": Si, "ray_start": rs, "ray_end": re, "box_warp": bw,
                "white_back": wb, "disparity_space_sampling": disp, "clamp_mode": "softplus"}
        g = torch.Generator().manual_seed(14)
        rand_coarse = torch.rand(1, R, S, 1, generator=g)
        u = torch.rand(R, Si, generator=g)
        ren = ImportanceRenderer()
        orig_rand_like = torch.rand_like
        torch.rand_like = lambda t, **kw: rand_coarse.clone()
        try:
            with replay_rng([u], []), torch.no_grad():
                ref = ren(planes, dec, ro, rd, opts)
        finally:
            torch.rand_like = orig_rand_like
        with torch.no_grad():
            mine = E.importance_render(planes, decp, ro, rd, opts, rand_coarse, u)
        for k, (a, b) in enumerate(zip(ref, mine)):
            same(a, b, f"{name} out{k}")
        # stage fixtures: field, marcher, importance sampling on the coarse pass
        with torch.no_grad():
            depths = E.sample_stratified(1, R, rs, re, S, rand_coarse, disp)
            coords = (ro.unsqueeze(-2) + depths * rd.unsqueeze(-2)).reshape(1, -1, 3)
            feats_ref = sample_from_planes(generate_planes(), planes, coords, padding_mode="zeros", box_warp=bw)
            out_ref = dec(feats_ref, None)
            rgb_o, sig_o = E.run_model(planes, decp, coords, bw)
            same(out_ref["rgb"], rgb_o, "field rgb"); same(out_ref["sigma"], sig_o, "field sigma")
            colors = rgb_o.reshape(1, R, S, 3); dens = sig_o.reshape(1, R, S, 1)
            m_ref = MipRayMarcher2()(colors, dens, depths, opts)
            m_o = E.mip_ray_march(colors, dens, depths, wb)
            for a, b in zip(m_ref, m_o):
                same(a, b, "march")
            with replay_rng([u], []):
                fine_ref = ren.sample_importance(depths, m_ref[2], Si)
            fine_o = E.sample_importance(depths, m_ref[2], Si, u)
            same(fine_ref, fine_o, "importance")
        frac_inside = float(((coords.abs() * (2 / bw)) <= 1).all(-1).float().mean())
        print(f"{name}: ok  rgb_final mean {float(ref[3].mean()):.4f}  opacity mean {float(ref[5].mean()):.4f} "
              f"inside-box {frac_inside:.2f}  depth range {float(ref[4].min()):.3f}..{float(ref[4].max()):.3f}")
        np.savez_compressed(
            os.path.join(GOLD, f"g10_eg3d_{name}.npz"),
            res=res, R=R, S=S, Si=Si, ray_start=rs, ray_end=re, box_warp=bw, white_back=wb, disparity=disp,
            planes_seed=11, dec_seed=12, rays_seed=13, rand_coarse=rand_coarse.numpy(), u=u.numpy(),
            depths_coarse=depths.numpy(), field_rgb=out_ref["rgb"].numpy(), field_sigma=out_ref["sigma"].numpy(),
            march_rgb=m_ref[0].numpy(), march_depth=m_ref[1].numpy(), march_weights=m_ref[2].numpy(),
            depths_fine=fine_ref.numpy(),
            rgb_coarse=ref[0].numpy(), depth_coarse=ref[1].numpy(), opacity_coarse=ref[2].numpy(),
            rgb_final=ref[3].numpy(), depth_final=ref[4].numpy(), opacity_final=ref[5].numpy())


if __name__ == "__main__":
    main()
