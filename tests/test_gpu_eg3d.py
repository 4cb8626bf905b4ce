"""GPU parity tests for kernel (6), the EG3D tri-plane renderer, through the C ABI, against the
reference-produced fixtures g10_eg3d_* and the CPU oracle.  Tolerances (fp32; the kernel's reduction orders differ from
MKL's / torch.sum's): field rgb 2e-5, sigma 1e-4 (|sigma| up to ~30); marcher 2e-5; end-to-end rgb / opacity 2e-4,
depth 1e-3 of a ~2.5 range; importance depths compared as sorted sets to 1e-3."""
import pytest
import torch

from oracle import eg3d_oracle as E
from tests._util import gold
from tests.test_eg3d_oracle_golden import CASES, setup

pytestmark = pytest.mark.gpu
DEV = "cuda"
KEYS = ["rgb_coarse", "depth_coarse", "opacity_coarse", "rgb_final", "depth_final", "opacity_final"]


def nb():
    import nerf_siren_b200
    return nerf_siren_b200


def dec_flat(dec):
    return torch.cat([dec["net.0.weight"].reshape(-1), dec["net.0.bias"], dec["net.2.weight"].reshape(-1), dec["net.2.bias"]]).to(DEV)


def test_relayout_is_a_pure_permutation():
    planes = torch.randn(3, 32, 24, 40, device=DEV)
    assert torch.equal(nb().ops.triplane_relayout(planes), planes.permute(0, 2, 3, 1).contiguous())


@pytest.fixture(params=["1", "0"], ids=["tensor-core-decoder", "cuda-core-decoder"])
def decoder_arm(request, monkeypatch):
    """NERFB_TP_MMA: 1 (default) = the decoder's first layer as mma.sync TF32 x3 on 16 samples at a time, 0 = one sample at a
    time on CUDA cores.  Read per call by the library; both arms carry the same tolerances."""
    monkeypatch.setenv("NERFB_TP_MMA", request.param)
    return request.param


@pytest.mark.parametrize("name", CASES)
def test_field_matches_reference(name, decoder_arm):
    g = gold("g10_eg3d_" + name)
    planes, dec, ro, rd, opts = setup(g)
    R, S = int(g["R"]), int(g["S"])
    coords = (ro.unsqueeze(-2) + g["depths_coarse"] * rd.unsqueeze(-2)).reshape(-1, 3)
    cl = nb().ops.triplane_relayout(planes[0].to(DEV))
    rgb, sigma = nb().ops.triplane_field(cl, dec_flat(dec), 1.0, coords.to(DEV).contiguous(), opts["box_warp"])
    torch.testing.assert_close(rgb.cpu(), g["field_rgb"][0], rtol=0, atol=2e-5)
    torch.testing.assert_close(sigma.cpu(), g["field_sigma"][0, :, 0], rtol=1e-5, atol=1e-4)


def test_field_outside_box_and_nonfinite_coordinates(decoder_arm):
    """zeros padding: outside the planes the features are 0 and the decoder sees only its biases; NaN coordinates tap nothing."""
    g = gold("g10_eg3d_small")
    planes, dec, _, _, opts = setup(g)
    coords = torch.tensor([[5.0, 5.0, 5.0], [-9.0, 0.0, 0.0], [float("nan"), 0.0, 0.0], [0.499999, -0.499999, 0.5], [0.0, 0.0, 0.0]])
    ref_rgb, ref_sigma = E.run_model(planes, dec, coords[None], 1.0)
    cl = nb().ops.triplane_relayout(planes[0].to(DEV))
    rgb, sigma = nb().ops.triplane_field(cl, dec_flat(dec), 1.0, coords.to(DEV), 1.0)
    ok = [0, 1, 3, 4]
    torch.testing.assert_close(rgb.cpu()[ok], ref_rgb[0][ok], rtol=0, atol=2e-5)
    torch.testing.assert_close(sigma.cpu()[ok], ref_sigma[0, ok, 0], rtol=1e-5, atol=1e-4)
    assert torch.isfinite(rgb[2]).all()               # x = NaN: planes (x,y), (x,z), (z,x) all untapped -> bias-only output
    torch.testing.assert_close(rgb[2].cpu(), ref_rgb[0][0], rtol=0, atol=2e-5)


@pytest.mark.parametrize("name", CASES)
def test_marcher_matches_reference(name):
    g = gold("g10_eg3d_" + name)
    R, S = int(g["R"]), int(g["S"])
    rgb, depth, w = nb().ops.mip_march(g["field_rgb"].reshape(R, S, 3).to(DEV), g["field_sigma"].reshape(R, S).to(DEV),
                                       g["depths_coarse"].reshape(R, S).to(DEV), bool(g["white_back"]))
    torch.testing.assert_close(w.cpu(), g["march_weights"].reshape(R, S - 1), rtol=1e-5, atol=2e-6)
    torch.testing.assert_close(rgb.cpu(), g["march_rgb"][0], rtol=0, atol=2e-5)
    torch.testing.assert_close(depth.cpu(), g["march_depth"][0, :, 0], rtol=0, atol=2e-5)


def test_marcher_zero_weight_rays_clamp_to_global_max_depth():
    """sum w = 0 -> 0/0 = NaN -> +inf -> clamped to max(depths) over the WHOLE call (ray_marcher.py:46-50)."""
    R, S = 5, 8
    depths = torch.linspace(1, 2, S).repeat(R, 1) + torch.arange(R)[:, None] * 0.1
    dens = torch.full((R, S), -200.0); dens[0] = 3.0                   # softplus(-201) == 0 exactly -> alpha = 0
    colors = torch.rand(R, S, 3)
    ref = E.mip_ray_march(colors[None], dens[None, ..., None], depths[None, ..., None])
    rgb, depth, w = nb().ops.mip_march(colors.to(DEV), dens.to(DEV), depths.to(DEV))
    torch.testing.assert_close(depth.cpu(), ref[1][0, :, 0], rtol=0, atol=1e-6)
    assert float(depth[1]) == float(depths.max())
    torch.testing.assert_close(rgb.cpu(), ref[0][0], rtol=0, atol=1e-6)


@pytest.mark.parametrize("name", CASES)
def test_fused_render_matches_reference(name, decoder_arm):
    g = gold("g10_eg3d_" + name)
    planes, dec, ro, rd, opts = setup(g)
    R, S, Si = int(g["R"]), int(g["S"]), int(g["Si"])
    cl = nb().ops.triplane_relayout(planes[0].to(DEV))
    out = nb().ops.triplane_render(cl, dec_flat(dec), 1.0, ro[0].to(DEV), rd[0].to(DEV), S, Si, opts["ray_start"], opts["ray_end"],
                                   None, None, opts["disparity_space_sampling"], opts["box_warp"], opts["white_back"],
                                   g["rand_coarse"].reshape(R, S).to(DEV), g["u"].to(DEV), 0, return_depths_fine=True)
    tol = {"rgb": 2e-4, "opa": 2e-4, "dep": 1e-3}
    for got, key in zip(out[:6], KEYS):
        torch.testing.assert_close(got.cpu().reshape(g[key][0].shape), g[key][0], rtol=0, atol=tol[key[:3]])
    fine = torch.sort(out[6].cpu(), -1)[0]
    ref_fine = torch.sort(g["depths_fine"].reshape(R, Si), -1)[0]
    torch.testing.assert_close(fine, ref_fine, rtol=0, atol=1e-3)
    assert float((fine - ref_fine).abs().mean()) < 2e-5


def test_renderer_module_is_a_drop_in():
    """ImportanceRenderer / OSGDecoder with the reference's constructor arguments, state-dict keys and 6-tuple output."""
    g = gold("g10_eg3d_small")
    planes, dec, ro, rd, opts = setup(g)
    m = nb().eg3d
    decoder = m.OSGDecoder(32, {"decoder_lr_mul": 1, "decoder_output_dim": 3}).to(DEV)
    assert list(decoder.state_dict().keys()) == ["net.0.weight", "net.0.bias", "net.2.weight", "net.2.bias"]
    decoder.load_state_dict(dec)
    ren = m.ImportanceRenderer()
    R, S = int(g["R"]), int(g["S"])
    with torch.no_grad():
        out = ren(planes.to(DEV), decoder, ro.to(DEV), rd.to(DEV), opts, rand_coarse=g["rand_coarse"].to(DEV), u=g["u"].to(DEV))
    assert [tuple(o.shape) for o in out] == [(1, R, 3), (1, R, 1), (1, R, 1), (1, R, 3), (1, R, 1), (1, R, 1)]
    for got, key in zip(out, KEYS):
        torch.testing.assert_close(got.cpu(), g[key], rtol=0, atol=1e-3 if "depth" in key else 2e-4)
    coords = (ro.unsqueeze(-2) + g["depths_coarse"] * rd.unsqueeze(-2)).reshape(1, -1, 3)
    with torch.no_grad():
        rm = ren.run_model(planes.to(DEV), decoder, coords.to(DEV), None, opts)
    torch.testing.assert_close(rm["rgb"].cpu(), g["field_rgb"], rtol=0, atol=2e-5)
    assert rm["sigma"].shape == g["field_sigma"].shape
    mm = m.MipRayMarcher2()(g["field_rgb"].reshape(1, R, S, 3).to(DEV), g["field_sigma"].reshape(1, R, S, 1).to(DEV),
                            g["depths_coarse"].to(DEV), opts)
    torch.testing.assert_close(mm[2].cpu(), g["march_weights"], rtol=1e-5, atol=2e-6)
    with torch.no_grad():                                  # internal RNG path: finite, reproducible per seed counter
        a = m.ImportanceRenderer()(planes.to(DEV), decoder, ro.to(DEV), rd.to(DEV), opts)
        b = m.ImportanceRenderer()(planes.to(DEV), decoder, ro.to(DEV), rd.to(DEV), opts)
    assert all(torch.equal(x, y) for x, y in zip(a, b)) and all(torch.isfinite(x).all() for x in a)


def test_full_size_properties_config5(decoder_arm):
    """BASELINE config 5 shape: 3 x 32 x 256^2 planes, 128^2 = 16,384 rays, 48 + 48 samples.  Size-independent
    properties: opacity in [0, 1], colours inside the sigmoid clamp range, depths inside [ray_start, ray_end + delta],
    white_back identity rgb_wb = rgb + 1 - opacity, and per-ray independence (a sub-batch renders identically)."""
    planes = E.planes_init(seed=5, res=256).to(DEV)
    dec = dec_flat(E.decoder_init(seed=6))
    ro, rd = E.camera_rays(16384, seed=7)
    ro, rd = ro[0].to(DEV), rd[0].to(DEV)
    g = torch.Generator().manual_seed(8)
    rc, u = torch.rand(16384, 48, generator=g).to(DEV), torch.rand(16384, 48, generator=g).to(DEV)
    cl = nb().ops.triplane_relayout(planes[0])
    a = nb().ops.triplane_render(cl, dec, 1.0, ro, rd, 48, 48, 2.2, 3.3, None, None, False, 1.0, False, rc, u, 0)
    w = nb().ops.triplane_render(cl, dec, 1.0, ro, rd, 48, 48, 2.2, 3.3, None, None, False, 1.0, True, rc, u, 0)
    for t in a:
        assert torch.isfinite(t).all()
    assert float(a[5].min()) >= 0 and float(a[5].max()) <= 1 + 1e-5
    assert float(a[3].min()) >= -0.001 - 1e-6 and float(a[3].max()) <= 1.001 + 1e-6
    assert float(a[4].min()) >= 2.2 and float(a[4].max()) <= 3.3 + 1.1 / 47 + 1e-6
    torch.testing.assert_close(w[3], a[3] + 1 - a[5][:, None], rtol=0, atol=2e-6)
    sub = nb().ops.triplane_render(cl, dec, 1.0, ro[:1000].contiguous(), rd[:1000].contiguous(), 48, 48, 2.2, 3.3, None, None, False,
                                   1.0, False, rc[:1000].contiguous(), u[:1000].contiguous(), 0)
    assert torch.equal(sub[3], a[3][:1000]) and torch.equal(sub[5], a[5][:1000])      # depth differs only through the global clamp


def test_empty_and_bad_arguments():
    cl = torch.zeros(3, 8, 8, 32, device=DEV)
    dec = torch.zeros(64 * 32 + 64 + 256 + 4, device=DEV)
    out = nb().ops.triplane_render(cl, dec, 1.0, torch.zeros(0, 3, device=DEV), torch.zeros(0, 3, device=DEV), 16, 16, 1.0, 2.0)
    assert out[0].shape == (0, 3)
    with pytest.raises(ValueError):
        nb().ops.triplane_relayout(torch.zeros(2, 32, 8, 8, device=DEV))
    with pytest.raises(RuntimeError):
        nb().ops.triplane_relayout(torch.zeros(3, 16, 8, 8, device=DEV))          # 16 channels: NERFB_E_UNSUPPORTED
    with pytest.raises(RuntimeError):
        nb().ops.triplane_field(torch.zeros(3, 8, 8, 32), dec, 1.0, torch.zeros(4, 3), 1.0)   # CPU tensors are refused


def test_fresh_plane_tensors_are_never_served_from_the_cache():
    """a plane source that returns a FRESH tensor per call (generator backbone) usually gets the address the
    caching allocator just freed, version 0, same shape.  Two different plane sets rendered back to back through one
    renderer must give two different images, each equal to a fresh renderer's; a persistent Parameter updated in place
    (version bump) must be re-laid out too."""
    g = gold("g10_eg3d_small")
    planes, dec, ro, rd, opts = setup(g)
    m = nb().eg3d
    decoder = m.OSGDecoder(32, {"decoder_lr_mul": 1, "decoder_output_dim": 3}).to(DEV)
    decoder.load_state_dict(dec)
    kw = dict(rand_coarse=g["rand_coarse"].to(DEV), u=g["u"].to(DEV))
    ren = m.ImportanceRenderer()
    outs = []
    with torch.no_grad():
        for scale in (1.0, -0.5):
            fresh = (planes * scale).to(DEV)                       # new tensor object; freed before the next one is made
            outs.append(ren(fresh, decoder, ro.to(DEV), rd.to(DEV), opts, **kw)[3].clone())
            want = m.ImportanceRenderer()(fresh, decoder, ro.to(DEV), rd.to(DEV), opts, **kw)[3]
            assert torch.equal(outs[-1], want), scale
            del fresh, want
        assert not torch.equal(outs[0], outs[1])
        p = torch.nn.Parameter(planes.clone().to(DEV))
        a = ren(p, decoder, ro.to(DEV), rd.to(DEV), opts, **kw)[3].clone()
        a2 = ren(p, decoder, ro.to(DEV), rd.to(DEV), opts, **kw)[3]
        assert torch.equal(a, a2)
        p.mul_(0.25)                                                # in-place update: same object, new _version
        b = ren(p, decoder, ro.to(DEV), rd.to(DEV), opts, **kw)[3]
        assert not torch.equal(a, b)


def test_untrainable_calls_take_the_fused_path():
    """Nothing requires grad -> the single fused launch (bit-identical with the no_grad call), grad mode or not.  The
    differentiable schedule is covered by tests/test_gpu_eg3d_train.py."""
    g = gold("g10_eg3d_small")
    planes, dec, ro, rd, opts = setup(g)
    m = nb().eg3d
    decoder = m.OSGDecoder(32, {"decoder_lr_mul": 1, "decoder_output_dim": 3}).to(DEV)
    for p in decoder.parameters():
        p.requires_grad_(False)
    kw = dict(rand_coarse=g["rand_coarse"].to(DEV), u=g["u"].to(DEV))
    ren = m.ImportanceRenderer()
    out = ren(planes.to(DEV), decoder, ro.to(DEV), rd.to(DEV), opts, **kw)
    with torch.no_grad():
        out2 = ren(planes.to(DEV), decoder, ro.to(DEV), rd.to(DEV), opts, **kw)
    assert all(torch.equal(a, b) for a, b in zip(out, out2)) and not out[3].requires_grad


