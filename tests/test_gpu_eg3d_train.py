"""GPU parity tests for the EG3D training backward (SURVEY rows a11-a13 widened to the reference's training step,
system.py:111-120): nerfb_triplane_field_bwd and nerfb_mip_march_bwd through the C ABI and through the host mirror's
autograd functions, against torch autograd over the CPU oracle (oracle/eg3d_oracle.py, fp32; it is pinned to the
reference by the g10_eg3d_* fixtures).  Tolerances (fp32 on both sides; the kernels accumulate plane gradients with
atomics and the decoder gradients in another order): relative L2 1e-4 for the unit tests, 2e-3 end to end (the fused
render's importance depths differ from the oracle's in the last bits and can land on the other side of a neighbouring
sample for a handful of rays; the reference detaches them, so the gradient is only piecewise smooth in them)."""
import pytest
import torch

from oracle import eg3d_oracle as E
from tests._util import gold
from tests.test_eg3d_oracle_golden import setup

pytestmark = pytest.mark.gpu
DEV = "cuda"
DEC_KEYS = ["net.0.weight", "net.0.bias", "net.2.weight", "net.2.bias"]


def nb():
    import nerf_siren_b200
    return nerf_siren_b200


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def flat(dec):
    return torch.cat([dec[k].reshape(-1) for k in DEC_KEYS])


def leaf_dec(seed, lr_mul=1.0):
    dec = E.decoder_init(seed=seed)
    if lr_mul != 1.0:                              # FullyConnectedLayer stores weight / lr_multiplier
        dec = {k: (v / lr_mul if k.endswith("weight") else v) for k, v in dec.items()}
    return {k: v.clone().requires_grad_(True) for k, v in dec.items()}


@pytest.fixture(params=["1", "0"], ids=["tensor-core-decoder", "cuda-core-decoder"])
def decoder_arm(request, monkeypatch):
    """NERFB_TP_MMA: 1 (default) = decoder GEMMs as mma.sync TF32 x3, 0 = the CUDA-core kernels.  Read per call by the library."""
    monkeypatch.setenv("NERFB_TP_MMA", request.param)
    return request.param


@pytest.mark.parametrize("res,n,lr_mul,box", [(16, 700, 1.0, 1.0), (64, 5000, 0.5, 2.0), (33, 257, 2.0, 1.0)])
def test_field_backward_matches_autograd(res, n, lr_mul, box, decoder_arm):
    """d(planes), d(decoder) of run_model for random cotangents; a third of the points lie outside the box (zero padding:
    no plane gradient from the taps that fall off the edge) and the plane is not square-friendly in the third case."""
    planes = (E.planes_init(seed=res, res=res)[0] * 0.7).requires_grad_(True)              # (3, 32, res, res)
    dec = leaf_dec(res + 1, lr_mul)
    g = torch.Generator().manual_seed(res + 2)
    coords = (torch.rand(n, 3, generator=g) - 0.5) * box * 1.5
    d_rgb, d_sigma = torch.randn(n, 3, generator=g), torch.randn(n, generator=g)
    rgb, sigma = E.run_model(planes[None], dec, coords[None], box, lr_mul)
    (rgb[0] * d_rgb).sum().add((sigma[0, :, 0] * d_sigma).sum()).backward()
    ops = nb().ops
    cl = ops.triplane_relayout(planes.detach().to(DEV))
    d_cl, d_dec = ops.triplane_field_bwd(cl, flat(dec).detach().to(DEV), lr_mul, coords.to(DEV), box, d_rgb.to(DEV), d_sigma.to(DEV))
    assert rel(d_cl.permute(0, 3, 1, 2).cpu(), planes.grad) < 1e-4
    assert rel(d_dec.cpu(), torch.cat([dec[k].grad.reshape(-1) for k in DEC_KEYS])) < 1e-4
    # determinism of the decoder part is not promised (atomics), but a second call must agree to rounding
    d_cl2, d_dec2 = ops.triplane_field_bwd(cl, flat(dec).detach().to(DEV), lr_mul, coords.to(DEV), box, d_rgb.to(DEV), d_sigma.to(DEV))
    assert rel(d_cl2, d_cl) < 1e-5 and rel(d_dec2, d_dec) < 1e-5


def test_field_backward_edges(decoder_arm):
    ops = nb().ops
    cl = torch.randn(3, 8, 8, 32, device=DEV)
    dec = torch.randn(2372, device=DEV) * 0.1
    d_cl, d_dec = ops.triplane_field_bwd(cl, dec, 1.0, torch.zeros(0, 3, device=DEV), 1.0, torch.zeros(0, 3, device=DEV), torch.zeros(0, device=DEV))
    assert d_cl.shape == cl.shape and not d_cl.any() and not d_dec.any()
    # all points outside the box: the decoder still sees its biases (gradients on net.0.bias / net.2.*), planes get none
    co = torch.full((64, 3), 7.0, device=DEV)
    d_cl, d_dec = ops.triplane_field_bwd(cl, dec, 1.0, co, 1.0, torch.ones(64, 3, device=DEV), torch.ones(64, device=DEV))
    assert not d_cl.any() and not d_dec[:2048].any() and d_dec[2048:].abs().sum() > 0
    # NaN coordinate: taps nothing, contributes like an outside point, poisons nothing
    co[0, 0] = float("nan")
    d_cl, d_dec = ops.triplane_field_bwd(cl, dec, 1.0, co, 1.0, torch.ones(64, 3, device=DEV), torch.ones(64, device=DEV))
    assert torch.isfinite(d_cl).all() and torch.isfinite(d_dec).all()
    with pytest.raises(RuntimeError):
        ops.triplane_field_bwd(cl.cpu(), dec, 1.0, co, 1.0, torch.ones(64, 3, device=DEV), torch.ones(64, device=DEV))


@pytest.mark.parametrize("R,S,white", [(37, 24, False), (200, 96, True), (5, 2, False), (64, 129, True)])
def test_marcher_backward_matches_autograd(R, S, white):
    """All three cotangents (rgb, depth, weights) at once; S = 2 is the single-interval edge, S = 129 crosses the 32-lane
    chunks of the scan; densities span dead (-20) to saturated (+30) intervals."""
    g = torch.Generator().manual_seed(R * 1000 + S)
    colors = torch.rand(1, R, S, 3, generator=g).requires_grad_(True)
    dens = (torch.randn(1, R, S, 1, generator=g) * 6 + 1).requires_grad_(True)
    depths = (2.0 + torch.rand(1, R, S, 1, generator=g).cumsum(2) * (1.0 / S))
    d_rgb, d_depth, d_w = torch.randn(R, 3, generator=g), torch.randn(R, generator=g), torch.randn(R, S - 1, generator=g)
    rgb, depth, w = E.mip_ray_march(colors, dens, depths, white)
    ((rgb[0] * d_rgb).sum() + (depth[0, :, 0] * d_depth).sum() + (w[0, :, :, 0] * d_w).sum()).backward()
    ops = nb().ops
    args = (colors.detach()[0].to(DEV).contiguous(), dens.detach()[0, :, :, 0].to(DEV).contiguous(), depths[0, :, :, 0].to(DEV).contiguous())
    d_c, d_s = ops.mip_march_bwd(*args, white, d_rgb.to(DEV), d_depth.to(DEV), d_w.to(DEV))
    assert rel(d_c.cpu(), colors.grad[0]) < 1e-4
    assert rel(d_s.cpu(), dens.grad[0, :, :, 0]) < 1e-4
    # each cotangent alone (the other pointers null)
    for k, kw in enumerate([dict(d_rgb=d_rgb.to(DEV)), dict(d_depth=d_depth.to(DEV)), dict(d_weights=d_w.to(DEV))]):
        colors.grad = dens.grad = None
        rgb, depth, w = E.mip_ray_march(colors, dens, depths, white)
        [(rgb[0] * d_rgb).sum(), (depth[0, :, 0] * d_depth).sum(), (w[0, :, :, 0] * d_w).sum()][k].backward()
        d_c, d_s = ops.mip_march_bwd(*args, white, **kw)
        assert rel(d_s.cpu(), dens.grad[0, :, :, 0]) < 1e-4, k
        if colors.grad is not None and colors.grad.abs().sum() > 0:
            assert rel(d_c.cpu(), colors.grad[0]) < 1e-4, k
        else:
            assert not d_c.any()


def test_marcher_backward_empty_ray_has_zero_depth_gradient():
    """A ray whose weights sum to 0 gets depth = nan -> inf -> clamp (ray_marcher.py:45-47): no gradient through it."""
    ops = nb().ops
    colors = torch.rand(3, 8, 3, device=DEV)
    dens = torch.full((3, 8), -1e4, device=DEV)
    depths = torch.linspace(1, 2, 8, device=DEV).expand(3, 8).contiguous()
    d_c, d_s = ops.mip_march_bwd(colors, dens, depths, False, None, torch.ones(3, device=DEV), None)
    assert torch.isfinite(d_s).all() and not d_s.any() and not d_c.any()


@pytest.mark.parametrize("name", ["small", "white", "disp"])
def test_training_step_gradients_match_oracle_autograd(name, decoder_arm):
    """The host mirror's differentiable schedule end to end (ImportanceRenderer.forward with trainable planes and decoder,
    the reference's loss: MSE of coarse + fine colours, system.py:111-120 / losses.py) against autograd over the oracle on
    the fixtures' inputs and random numbers."""
    gd = gold("g10_eg3d_" + name)
    planes, dec0, ro, rd, opts = setup(gd)
    target = torch.rand(ro.shape[1], 3, generator=torch.Generator().manual_seed(3))
    # oracle
    p_ref = planes.clone().requires_grad_(True)
    dec = {k: v.clone().requires_grad_(True) for k, v in dec0.items()}
    out = E.importance_render(p_ref, dec, ro, rd, opts, gd["rand_coarse"], gd["u"])
    loss_ref = (((out[0][0] - target) ** 2).mean() + ((out[3][0] - target) ** 2).mean() + 0.1 * out[5].mean() + 0.05 * out[4].mean()
                + 0.07 * out[2].mean() + 0.03 * out[1].mean())
    loss_ref.backward()
    # this package
    m = nb().eg3d
    decoder = m.OSGDecoder(32, {"decoder_lr_mul": 1, "decoder_output_dim": 3})
    decoder.load_state_dict(dec0)
    decoder = decoder.to(DEV)
    p = torch.nn.Parameter(planes.clone().to(DEV))
    got = m.ImportanceRenderer()(p, decoder, ro.to(DEV), rd.to(DEV), opts, rand_coarse=gd["rand_coarse"].to(DEV), u=gd["u"].to(DEV))
    for a, b in zip(got, out):
        torch.testing.assert_close(a.detach().cpu(), b.detach(), rtol=0, atol=1e-3)
    with torch.no_grad():                                   # the training forward IS the fused render launch: identical outputs
        same = m.ImportanceRenderer()(p, decoder, ro.to(DEV), rd.to(DEV), opts, rand_coarse=gd["rand_coarse"].to(DEV), u=gd["u"].to(DEV))
    assert all(torch.equal(a.detach(), b) for a, b in zip(got, same)) and got[3].requires_grad
    t = target.to(DEV)
    loss = (((got[0][0] - t) ** 2).mean() + ((got[3][0] - t) ** 2).mean() + 0.1 * got[5].mean() + 0.05 * got[4].mean()
            + 0.07 * got[2].mean() + 0.03 * got[1].mean())
    assert abs(float(loss.detach()) - float(loss_ref.detach())) < 1e-4
    loss.backward()
    assert rel(p.grad.cpu(), p_ref.grad) < 2e-3
    sd = dict(decoder.named_parameters())
    for k in DEC_KEYS:
        assert rel(sd[k].grad.cpu(), dec[k].grad) < 2e-3, k


def test_training_reduces_the_loss_and_eval_path_sees_the_update():
    """Ten Adam steps on planes + decoder through EG3D_Renderer.render (the object NeRFSystem_eg3d trains,
    eg3d_renderer.py:6-67): the loss drops, and the fused no_grad render reflects the updated planes (cache keyed on
    _version)."""
    m = nb().eg3d
    torch.manual_seed(0)
    kw = dict(m.DEFAULT_RENDERING_KWARGS, depth_resolution=24, depth_resolution_importance=24, ray_start=2.2, ray_end=3.3, box_warp=1.0)
    ren = m.EG3D_Renderer(m.LearnedPlanes(resolution=64), kw).to(DEV)
    ro, rd = E.camera_rays(512, seed=1)
    ro, rd = ro[0].to(DEV), rd[0].to(DEV)
    target = torch.rand(512, 3, device=DEV) * 0.5 + 0.25
    opt = torch.optim.Adam(ren.parameters(), lr=2e-2)
    losses = []
    with torch.no_grad():
        before = ren.render(-1, ro, rd)["rgb_fine"].clone()
    for _ in range(10):
        opt.zero_grad()
        out = ren.render(-1, ro, rd)
        loss = ((out["rgb_coarse"] - target) ** 2).mean() + ((out["rgb_fine"] - target) ** 2).mean()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    assert losses[-1] < 0.7 * losses[0], losses
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in ren.parameters())
    with torch.no_grad():
        after = ren.render(-1, ro, rd)["rgb_fine"]
    assert float(((after - target) ** 2).mean()) < float(((before - target) ** 2).mean())


def test_render_train_entry_writes_the_samples_of_the_render():
    """nerfb_triplane_render_train = nerfb_triplane_render + samples (R, 5, S + Si): same six outputs bit for bit; the sample
    depths are the stratified then the importance depths; sigma / colour equal the stand-alone field at those points; and
    ray_start / ray_end arrays plus the in-kernel generator are accepted."""
    gd = gold("g10_eg3d_small")
    planes, dec0, ro, rd, opts = setup(gd)
    ops = nb().ops
    R, S, Si = int(gd["R"]), int(gd["S"]), int(gd["Si"])
    cl = ops.triplane_relayout(planes[0].to(DEV))
    dec = flat(dec0).to(DEV)
    args = (cl, dec, 1.0, ro[0].to(DEV), rd[0].to(DEV), S, Si, opts["ray_start"], opts["ray_end"], None, None, False, opts["box_warp"], False,
            gd["rand_coarse"].reshape(R, S).to(DEV), gd["u"].reshape(R, Si).to(DEV), 0)
    plain = ops.triplane_render(*args, return_depths_fine=True)
    train = ops.triplane_render(*args, return_samples=True)
    assert all(torch.equal(a, b) for a, b in zip(plain[:6], train[:6]))
    smp = train[6]
    assert smp.shape == (R, 5, S + Si) and torch.equal(smp[:, 0, S:], plain[6])
    torch.testing.assert_close(smp[:, 0, :S].cpu(), gd["depths_coarse"].reshape(R, S), rtol=0, atol=1e-6)
    coords = (ro[0].to(DEV).unsqueeze(1) + smp[:, 0].unsqueeze(-1) * rd[0].to(DEV).unsqueeze(1)).reshape(-1, 3).contiguous()
    rgb, sigma = ops.triplane_field(cl, dec, 1.0, coords, opts["box_warp"])
    torch.testing.assert_close(smp[:, 1].reshape(-1), sigma, rtol=1e-5, atol=1e-4)
    torch.testing.assert_close(smp[:, 2:5].transpose(1, 2).reshape(-1, 3), rgb, rtol=0, atol=2e-5)
    lo, hi = torch.full((R,), 2.3, device=DEV), torch.full((R,), 3.2, device=DEV)
    auto = ops.triplane_render(cl, dec, 1.0, ro[0].to(DEV), rd[0].to(DEV), S, Si, 0.0, 0.0, lo, hi, False, opts["box_warp"], True, None, None, 7,
                               return_samples=True)
    assert torch.isfinite(auto[6]).all() and float(auto[6][:, 0].min()) >= 2.3
    with pytest.raises(ValueError):
        ops.triplane_render(*args, return_depths_fine=True, return_samples=True)


@pytest.mark.parametrize("R,S,Si,white,unsorted", [(300, 24, 24, False, False), (64, 48, 48, True, False), (33, 12, 40, False, False),
                                                   (40, 48, 48, False, True)])
def test_samples_backward_equals_the_marcher_backward_composition(R, S, Si, white, unsorted):
    """nerfb_triplane_samples_bwd against the same thing composed from nerfb_mip_march_bwd and torch's stable sort / gather /
    scatter (renderer.py:160-170), all six cotangents at once and the importance depths interleaved with (and in 12 + 40
    outnumbering) the coarse ones; duplicates among the depths exercise the (value, index) tie rule."""
    ops = nb().ops
    g = torch.Generator(device=DEV).manual_seed(R + S)
    St = S + Si
    coarse = 2.0 + torch.rand(R, S, device=DEV, generator=g).cumsum(1) / S
    if unsorted:                                         # the kernels take the merge shortcut only for ascending coarse depths:
        coarse[1::2] = coarse[1::2].flip(1)              # every other ray descending -> the full (value, index) count
    fine = 2.0 + torch.rand(R, Si, device=DEV, generator=g) * 1.1
    fine[:, 0] = coarse[:, S // 2]                       # a tie between a coarse and an importance depth
    fine[:, 1] = fine[:, 2]                              # and between two importance depths
    samples = torch.empty(R, 5, St, device=DEV)
    samples[:, 0] = torch.cat([coarse, fine], 1)
    samples[:, 1] = torch.randn(R, St, device=DEV, generator=g) * 5 + 1
    samples[:, 2:5] = torch.rand(R, 3, St, device=DEV, generator=g)
    o, d = torch.randn(R, 3, device=DEV, generator=g), torch.randn(R, 3, device=DEV, generator=g)
    cot = [torch.randn(R, 3, device=DEV, generator=g), torch.randn(R, device=DEV, generator=g), torch.randn(R, device=DEV, generator=g),
           torch.randn(R, 3, device=DEV, generator=g), torch.randn(R, device=DEV, generator=g), torch.randn(R, device=DEV, generator=g)]
    coords, d_rgb, d_sig = ops.triplane_samples_bwd(samples, o, d, S, Si, white, *cot)
    depth, sigma, rgb = samples[:, 0], samples[:, 1], samples[:, 2:5].transpose(1, 2)
    want_rgb, want_sig = torch.zeros(R, St, 3, device=DEV), torch.zeros(R, St, device=DEV)
    dc, ds = ops.mip_march_bwd(rgb[:, :S].contiguous(), sigma[:, :S].contiguous(), depth[:, :S].contiguous(), white, cot[0], cot[1],
                               cot[2].reshape(R, 1).expand(R, S - 1).contiguous())
    want_rgb[:, :S], want_sig[:, :S] = dc, ds
    all_d, idx = torch.sort(depth, dim=1, stable=True)
    idx3 = idx.unsqueeze(-1).expand(-1, -1, 3)
    dc, ds = ops.mip_march_bwd(torch.gather(rgb, 1, idx3).contiguous(), torch.gather(sigma, 1, idx), all_d, white, cot[3], cot[4],
                               cot[5].reshape(R, 1).expand(R, St - 1).contiguous())
    want_rgb.scatter_add_(1, idx3, dc)
    want_sig.scatter_add_(1, idx, ds)
    assert rel(d_rgb.view(R, St, 3), want_rgb) < 1e-6 and rel(d_sig.view(R, St), want_sig) < 1e-6
    torch.testing.assert_close(coords.view(R, St, 3), o.unsqueeze(1) + depth.unsqueeze(-1) * d.unsqueeze(1), rtol=0, atol=1e-6)
    # null cotangents = zero cotangents
    only_f = ops.triplane_samples_bwd(samples, o, d, S, Si, white, None, None, None, cot[3], None, None)
    zeros = [torch.zeros_like(c) for c in cot]
    zeros[3] = cot[3]
    ref_f = ops.triplane_samples_bwd(samples, o, d, S, Si, white, *zeros)
    assert torch.equal(only_f[1], ref_f[1]) and torch.equal(only_f[2], ref_f[2])


def test_run_model_is_differentiable_through_the_mirror():
    """renderer.run_model (what extract_color_mesh_eg3d.py samples, and what a density regulariser would differentiate)."""
    m = nb().eg3d
    planes = E.planes_init(seed=9, res=32)
    dec0 = E.decoder_init(seed=10)
    coords = (torch.rand(1, 300, 3, generator=torch.Generator().manual_seed(11)) - 0.5)
    p_ref = planes.clone().requires_grad_(True)
    rgb, sigma = E.run_model(p_ref, dec0, coords, 1.0)
    (rgb.sum() + (sigma ** 2).sum()).backward()
    decoder = m.OSGDecoder(32, {"decoder_lr_mul": 1, "decoder_output_dim": 3})
    decoder.load_state_dict(dec0)
    decoder = decoder.to(DEV)
    p = torch.nn.Parameter(planes.clone().to(DEV))
    out = m.ImportanceRenderer().run_model(p, decoder, coords.to(DEV), None, {"box_warp": 1.0})
    (out["rgb"].sum() + (out["sigma"] ** 2).sum()).backward()
    assert rel(p.grad.cpu(), p_ref.grad) < 1e-4
