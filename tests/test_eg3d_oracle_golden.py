"""oracle/eg3d_oracle.py against the committed REFERENCE outputs (tests/golden/g10_eg3d_*.npz, written by
oracle/make_goldens_eg3d.py from volumetric_rendering/ + eg3d_training/triplane.py imported read-only)."""
import pytest
import torch

from oracle import eg3d_oracle as E
from tests._util import gold

torch.set_num_threads(2)
CASES = ["small", "white", "disp", "cfg5"]


def setup(g):
    planes = E.planes_init(seed=int(g["planes_seed"]), res=int(g["res"]))
    dec = E.decoder_init(seed=int(g["dec_seed"]))
    ro, rd = E.camera_rays(int(g["R"]), seed=int(g["rays_seed"]))
    opts = {"depth_resolution": int(g["S"]), "depth_resolution_importance": int(g["Si"]), "ray_start": float(g["ray_start"]),
            "ray_end": float(g["ray_end"]), "box_warp": float(g["box_warp"]), "white_back": bool(g["white_back"]),
            "disparity_space_sampling": bool(g["disparity"]), "clamp_mode": "softplus"}
    return planes, dec, ro, rd, opts


@pytest.mark.parametrize("name", CASES)
def test_importance_render_matches_reference(name):
    g = gold("g10_eg3d_" + name)
    planes, dec, ro, rd, opts = setup(g)
    out = E.importance_render(planes, dec, ro, rd, opts, g["rand_coarse"], g["u"])
    for got, key in zip(out, ["rgb_coarse", "depth_coarse", "opacity_coarse", "rgb_final", "depth_final", "opacity_final"]):
        torch.testing.assert_close(got, g[key], rtol=0, atol=2e-6)          # bit-exact where it was generated; BLAS order elsewhere


@pytest.mark.parametrize("name", CASES[:3])
def test_stages_match_reference(name):
    g = gold("g10_eg3d_" + name)
    planes, dec, ro, rd, opts = setup(g)
    R, S = int(g["R"]), int(g["S"])
    depths = E.sample_stratified(1, R, opts["ray_start"], opts["ray_end"], S, g["rand_coarse"], opts["disparity_space_sampling"])
    assert torch.equal(depths, g["depths_coarse"])
    coords = (ro.unsqueeze(-2) + depths * rd.unsqueeze(-2)).reshape(1, -1, 3)
    rgb, sigma = E.run_model(planes, dec, coords, opts["box_warp"])
    torch.testing.assert_close(rgb, g["field_rgb"], rtol=0, atol=2e-6)
    torch.testing.assert_close(sigma, g["field_sigma"], rtol=0, atol=1e-5)
    m = E.mip_ray_march(g["field_rgb"].reshape(1, R, S, 3), g["field_sigma"].reshape(1, R, S, 1), depths, opts["white_back"])
    for got, key in zip(m, ["march_rgb", "march_depth", "march_weights"]):
        assert torch.equal(got, g[key])
    fine = E.sample_importance(depths, g["march_weights"], int(g["Si"]), g["u"])
    assert torch.equal(fine, g["depths_fine"])


def test_plane_projection_quirk():
    """planes 1 and 2 project to (x, z) and (z, x) (renderer.py:29-39) -- the upstream quirk the kernel hard-codes."""
    c = torch.tensor([[[0.1, 0.2, 0.3]]])
    p = E.project_onto_planes(E.generate_planes(), c)
    assert torch.allclose(p[:, 0], torch.tensor([[0.1, 0.2], [0.1, 0.3], [0.3, 0.1]]))
