"""GPU parity tests for kernels (2)(3)(4) and the stand-alone encoding: CUDA path through the C ABI vs the CPU
oracle / the reference-produced fixtures.  Bit-exact where the arithmetic is order-free (sampling, encoding
values, search indices), fp32 tolerances written per test elsewhere."""
import numpy as np
import pytest
import torch

from oracle import render_oracle as O
from tests._util import gold

pytestmark = pytest.mark.gpu
DEV = "cuda"


def nb():
    import nerf_siren_b200
    return nerf_siren_b200


@pytest.mark.parametrize("use_disp", [False, True])
@pytest.mark.parametrize("perturb", [0.0, 1.0, 0.5])
@pytest.mark.parametrize("S", [64, 48, 3, 1])
def test_sample_coarse_bit_exact(use_disp, perturb, S):
    for rays in (O.blender_rays(257, seed=1), O.llff_rays(129, seed=2)):
        if use_disp:
            rays = rays.clone(); rays[:, 6] = rays[:, 6].clamp_min(0.05)       # disparity needs near > 0
        pr = torch.rand(rays.shape[0], S, generator=torch.Generator().manual_seed(3))
        ref = O.stratified_z(rays, S, use_disp, perturb, pr)
        got = nb().ops.sample_coarse(rays.to(DEV), S, use_disp, perturb, pr.to(DEV) if perturb > 0 else None).cpu()
        assert torch.equal(got, ref)


def test_sample_coarse_empty_and_bad_args():
    z = nb().ops.sample_coarse(torch.zeros(0, 8, device=DEV), 64)
    assert z.shape == (0, 64)
    with pytest.raises(ValueError):
        nb().ops.sample_coarse(torch.zeros(4, 8, device=DEV), 64, perturb=1.0, perturb_rand=None)


def test_embedding_matches_reference_fixture():
    g = gold("g1_embed")
    e10 = nb().Embedding(3, 10)(g["x"].to(DEV)).cpu()
    e4 = nb().Embedding(3, 4)(g["d"].to(DEV)).cpu()
    assert e10.shape == (96, 63) and e4.shape == (96, 27)
    assert torch.equal(e10[:, :3], g["x"])                       # identity channels are exact
    # sin/cos of arguments up to 512*6 rad: CUDA sincosf vs the CPU libm differ by a few ulp of the result
    assert (e10 - g["e10"]).abs().max().item() <= 4e-7 * 1.0 + 2e-7
    assert (e4 - g["e4"]).abs().max().item() <= 4e-7
    assert torch.equal(nb().Embedding(3, 0)(g["x"].to(DEV)).cpu(), g["x"])


@pytest.mark.parametrize("S", [64, 128, 192])
@pytest.mark.parametrize("wb", [0, 1])
def test_composite_forward_backward(S, wb):
    g = gold("g6_composite")
    sig, rgbs, z, d, noise = (g[f"s{S}_{k}"] for k in ("sig", "rgbs", "z", "d", "noise"))
    n = sig.shape[0]
    rays = torch.zeros(n, 8); rays[:, 3:6] = d
    raw = torch.cat([rgbs, sig.unsqueeze(-1)], -1).contiguous()
    ops = nb().ops
    w, rgb, depth, op, _ = ops.composite_fwd(raw.to(DEV), z.to(DEV), rays.to(DEV), noise.to(DEV), 0.7, bool(wb))
    p = f"s{S}_wb{wb}_"
    # fp32, scan instead of sequential cumprod: tolerance 2e-6 relative to O(1) outputs
    torch.testing.assert_close(w.cpu(), g[p + "w"], rtol=2e-5, atol=2e-6)
    torch.testing.assert_close(rgb.cpu(), g[p + "rgb"], rtol=2e-5, atol=2e-6)
    torch.testing.assert_close(depth.cpu(), g[p + "depth"], rtol=2e-5, atol=1e-5)
    torch.testing.assert_close(op.cpu(), g[p + "w"].sum(1), rtol=2e-5, atol=2e-6)
    d_raw = ops.composite_bwd(raw.to(DEV), z.to(DEV), rays.to(DEV), noise.to(DEV), 0.7, bool(wb),
                              d_rgb=g[p + "up_rgb"].to(DEV), d_depth=g[p + "up_d"].to(DEV),
                              d_weights=g[p + "up_w"].to(DEV)).cpu()
    scale = g[p + "gsig"].abs().max().item()
    assert (d_raw[..., 3] - g[p + "gsig"]).abs().max().item() <= 2e-5 * scale + 1e-6
    torch.testing.assert_close(d_raw[..., :3], g[p + "grgb"], rtol=2e-5, atol=2e-6)


def test_composite_sigma_only_and_class_channels():
    torch.manual_seed(0)
    n, S, C = 37, 64, 10
    raw = torch.rand(n, S, C); raw[..., 3] = torch.randn(n, S) * 3
    z = torch.sort(torch.rand(n, S) * 4 + 2, -1)[0]
    rays = O.blender_rays(n, seed=4)
    ops = nb().ops
    ref = O.composite(raw[..., 3], z, rays[:, 3:6], None, 0.0, raw[..., :3], False, raw[..., 4:])
    w, rgb, depth, op, extra = ops.composite_fwd(raw.to(DEV), z.to(DEV), rays.to(DEV), None, 0.0, False)
    torch.testing.assert_close(extra.cpu(), ref[3], rtol=2e-5, atol=2e-6)
    torch.testing.assert_close(rgb.cpu(), ref[0], rtol=2e-5, atol=2e-6)
    w1, r1, d1, op1, e1 = ops.composite_fwd(raw[..., 3:4].contiguous().to(DEV), z.to(DEV), rays.to(DEV), None, 0.0, False)
    assert r1 is None and d1 is None and e1 is None
    torch.testing.assert_close(w1.cpu(), ref[2], rtol=2e-5, atol=2e-6)
    # backward with class channels against autograd of the oracle
    rawg = raw.clone().requires_grad_(True)
    o = O.composite(rawg[..., 3], z, rays[:, 3:6], None, 0.0, rawg[..., :3], True, rawg[..., 4:])
    up = [torch.randn_like(t) for t in (o[0], o[1], o[3])]
    (o[0] * up[0]).sum().add((o[1] * up[1]).sum()).add((o[3] * up[2]).sum()).backward()
    d_raw = ops.composite_bwd(raw.to(DEV), z.to(DEV), rays.to(DEV), None, 0.0, True, d_rgb=up[0].to(DEV),
                              d_depth=up[1].to(DEV), d_extra=up[2].to(DEV)).cpu()
    torch.testing.assert_close(d_raw, rawg.grad, rtol=1e-4, atol=2e-5 * rawg.grad.abs().max().item())


@pytest.mark.parametrize("side", ["left", "right"])
def test_searchsorted_bit_exact_reference_test_grid(side):
    """Mirrors torchsearchsorted/test/test_searchsorted.py:26-44 (sizes, both sides, 1-row broadcast)."""
    g = gold("g8_searchsorted")
    out = nb().ops.searchsorted(g["a"].to(DEV), g["v"].to(DEV), side=side)
    assert out.dtype == torch.long and torch.equal(out.cpu(), g[side])
    rs = np.random.RandomState(0)
    for Ba, Bv, A, V in [(1, 1, 1, 1), (1, 100, 50, 12), (100, 1, 500, 120), (200, 200, 1, 12), (100, 100, 500, 1),
                         (200, 200, 50, 120)]:
        for _ in range(5):
            a = np.sort(rs.rand(Ba, A).astype(np.float32), 1)
            v = rs.rand(Bv, V).astype(np.float32)
            v[:, :1] = a[:1, :1]
            got = nb().ops.searchsorted(torch.from_numpy(a).to(DEV), torch.from_numpy(v).to(DEV), side=side).cpu().numpy()
            assert np.array_equal(got, O.searchsorted_rows(a, v, side))
    pre = torch.empty(50, 100, dtype=torch.long, device=DEV)
    assert nb().ops.searchsorted(g["a"].to(DEV), g["v"].to(DEV), side=side, out=pre) is pre
    with pytest.raises(ValueError):
        nb().ops.searchsorted(torch.zeros(3, 4, device=DEV), torch.zeros(2, 4, device=DEV))


@pytest.mark.parametrize("dtype", ["float64", "int64", "int32", "int16", "int8", "uint8"])
@pytest.mark.parametrize("side", ["left", "right"])
def test_searchsorted_every_dispatched_dtype(dtype, side):
    """The reference kernel is AT_DISPATCH_ALL_TYPES (searchsorted_cuda_kernel.cu:134): the same grid of the reference's own
    test (Ba, Bv, A, V incl. 1-row broadcast, test_searchsorted.py:26-44) for every other scalar type, against numpy.
    Integer inputs are drawn from a small range so that ties exercise left / right."""
    rs = np.random.RandomState(3)
    npd = np.dtype(dtype)
    for Ba, Bv, A, V in [(1, 1, 1, 1), (1, 100, 50, 12), (100, 1, 500, 120), (200, 200, 1, 12), (100, 100, 500, 1),
                         (200, 200, 50, 120)]:
        if npd.kind == "f":
            a, v = rs.rand(Ba, A).astype(npd), rs.rand(Bv, V).astype(npd)
        else:
            lo, hi = (0, 100) if npd.kind == "u" else (-60, 60)
            a, v = rs.randint(lo, hi, (Ba, A)).astype(npd), rs.randint(lo, hi, (Bv, V)).astype(npd)
        a = np.sort(a, 1)
        v[:, :1] = a[:1, :1]
        got = nb().ops.searchsorted(torch.from_numpy(a).to(DEV), torch.from_numpy(v).to(DEV), side=side)
        assert got.dtype == torch.long and np.array_equal(got.cpu().numpy(), O.searchsorted_rows(a, v, side)), (Ba, Bv, A, V)
    with pytest.raises(ValueError):
        nb().ops.searchsorted(torch.zeros(2, 4, device=DEV), torch.zeros(2, 4, device=DEV, dtype=torch.float64))


@pytest.mark.parametrize("name", ["peaky", "flat", "zero", "rand"])
@pytest.mark.parametrize("ni", [64, 128])
def test_resampling_indices_bit_exact_on_identical_cdf(name, ni):
    """north_star: indices bit-exact against the reference for identical CDFs (fixtures come from the
    reference's own ops: cdf, u -> inds)."""
    g = gold("g7_sample_pdf")
    cdf = g[f"{name}_{ni}_cdf"].to(DEV)
    for u, want in ((g[f"{name}_{ni}_u"], g[f"{name}_{ni}_inds_rnd"]),
                    (O.linspace01(ni).expand(cdf.shape[0], ni).contiguous(), g[f"{name}_{ni}_inds_det"])):
        got = nb().ops.searchsorted(cdf, u.to(DEV), side="right").cpu()
        assert torch.equal(got, want)


def _pdf_tolerance(bins, weights, u, eps=1e-5):
    """Per-sample tolerance for comparing two fp32 evaluations of models/rendering.py:36-66 whose cdf differ by
    summation order (<= 2e-6).  The formula b0 + (u-c0)/denom*(b1-b0) amplifies a cdf perturbation dc by
    (b1-b0)/denom, and the reference's `denom < eps -> 1` switch (:63) is discontinuous: samples whose denom is
    within 3e-6 of eps may legitimately land anywhere in their bin (`stable` False)."""
    _, cdf, uu, inds = O.sample_pdf(bins, weights, u.shape[1], det=False, u=u, return_aux=True)
    below, above = (inds - 1).clamp_min(0), inds.clamp_max(weights.shape[1])
    c0, c1 = cdf.gather(1, below), cdf.gather(1, above)
    b0, b1 = bins.gather(1, below), bins.gather(1, above)
    denom = c1 - c0
    stable = (denom - eps).abs() > 3e-6
    # an index flip (u within the cdf noise of a knot) is harmless unless a bin touching that knot is "tiny"
    # (denom < eps -> 1 collapses the whole bin onto its lower edge, so the two sides of the knot disagree)
    dn = cdf[:, 1:] - cdf[:, :-1]
    tiny = torch.nn.functional.pad((dn < eps + 3e-6), (1, 1), value=False)     # tiny[:, j+1] <-> bin j
    for knot, ck in ((below, c0), (above, c1)):
        near = (uu - ck).abs() < 4e-6
        exact0 = (knot == 0) & (ck == 0)                                        # cdf[0] = 0 in both: cannot flip
        touch = tiny.gather(1, knot) | tiny.gather(1, knot + 1)                 # bins knot-1 and knot
        stable &= ~(near & touch & ~exact0)
    amp = torch.where(denom < eps, torch.ones_like(denom), denom)
    tol = 4e-6 + (b1 - b0).abs() * (4e-6 / amp).clamp_max(1.0)
    # u >= cdf[-1] (the det endpoint u = 1): the reference returns the last edge exactly; an implementation whose
    # cdf[-1] rounds above 1 interpolates inside the last bin instead -- same point up to cdf noise / last denom
    top = inds == cdf.shape[1]
    last_dn = torch.where(dn[:, -1:] < eps, torch.ones_like(dn[:, -1:]), dn[:, -1:])
    tol_top = 4e-6 + (bins[:, -1:] - bins[:, -2:-1]).abs() * (4e-6 / last_dn).clamp_max(1.0)
    tol = torch.where(top, tol_top.expand_as(tol), tol)
    stable &= ~(top & tiny[:, -2:-1].expand_as(top))
    return tol, stable, b0, b1


@pytest.mark.parametrize("name", ["peaky", "flat", "zero", "rand"])
@pytest.mark.parametrize("ni", [64, 128])
def test_sample_pdf_values_and_self_consistent_indices(name, ni):
    g = gold("g7_sample_pdf")
    bins, w = g["bins"], g[f"{name}_w"]
    ops = nb().ops
    u_det = O.linspace01(ni).expand(bins.shape[0], ni).contiguous()
    s_det, cdf, inds = ops.sample_pdf(bins.to(DEV), w.to(DEV), ni, det=True, return_aux=True)
    s_rnd = ops.sample_pdf(bins.to(DEV), w.to(DEV), ni, det=False, u=g[f"{name}_{ni}_u"].to(DEV))
    # the in-kernel cdf is a warp scan (the reference: sequential cumsum): <= 2e-6 apart
    assert (cdf.cpu() - g[f"{name}_{ni}_cdf"]).abs().max().item() <= 2e-6
    for got, ref, u in ((s_det.cpu(), g[f"{name}_{ni}_det"], u_det), (s_rnd.cpu(), g[f"{name}_{ni}_rnd"], g[f"{name}_{ni}_u"])):
        tol, stable, b0, b1 = _pdf_tolerance(bins, w, u)
        assert bool(((got - ref).abs() <= tol)[stable].all())
        assert stable.float().mean().item() > 0.9
        assert bool((got >= bins[:, :1] - 1e-6).all()) and bool((got <= bins[:, -1:] + 1e-6).all())
    # indices are exact with respect to the kernel's own cdf
    assert torch.equal(inds.cpu(), torch.searchsorted(cdf.cpu(), u_det, right=True))
    assert nb().sample_pdf(bins.to(DEV), w.to(DEV), ni, det=True).shape == (bins.shape[0], ni)


@pytest.mark.parametrize("det", [True, False])
@pytest.mark.parametrize("si", [64, 128])
def test_resample_fused_call_site(det, si):
    torch.manual_seed(1)
    n, sc = 130, 64
    rays = O.blender_rays(n, seed=5)
    z = O.stratified_z(rays, sc, False, 0.0 if det else 1.0, torch.rand(n, sc))
    w = torch.rand(n, sc) ** 6
    w[:7] = 0                                                     # zero-weight rows
    u = O.linspace01(si).expand(n, si).contiguous() if det else torch.rand(n, si)
    mid = 0.5 * (z[:, :-1] + z[:, 1:])
    z_new = O.sample_pdf(mid, w[:, 1:-1], si, det=det, u=u)
    ref = torch.sort(torch.cat([z, z_new], -1), -1)[0]
    got = nb().ops.resample(z.to(DEV), w.to(DEV), si, det, None if det else u.to(DEV)).cpu()
    assert got.shape == (n, sc + si)
    assert bool((got[:, 1:] >= got[:, :-1]).all())                # sortedness, exact
    assert bool((got >= z[:, :1]).all()) and bool((got <= z[:, -1:]).all())
    # sorting is 1-Lipschitz in the sup norm: per ray, |sorted got - sorted ref| <= the largest per-sample tolerance;
    # rays holding a sample on the reference's discontinuous `denom < eps` switch are only checked as above
    tol, stable, _, _ = _pdf_tolerance(mid, w[:, 1:-1], u)
    ok_rays = stable.all(1)
    assert ok_rays.float().mean().item() > 0.5
    err = (got - ref).abs().max(1)[0]
    assert bool((err <= tol.max(1)[0])[ok_rays].all())
    # every coarse depth survives the merge exactly (values-only sort of the concatenation)
    for r in (0, 7, n - 1):
        assert set(z[r].tolist()) <= set(got[r].tolist())
