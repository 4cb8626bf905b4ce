"""the steps either side of render_rays.  CPU: oracle vs the reference-produced fixture g11_io, the
Lightning-free system's checkpoint layout / schedulers.  GPU: kernels through the C ABI against the oracle."""
import os
import types

import numpy as np
import pytest
import torch

from oracle import io_oracle as IO
from oracle import render_oracle as O
from tests._util import gold

DEV = "cuda"


def hp(**kw):
    base = dict(N_samples=16, N_importance=16, use_disp=False, perturb=1.0, noise_std=1.0, chunk=32768, batch_size=256,
                lr=5e-4, optimizer="adam", lr_scheduler="steplr", decay_step=[2], decay_gamma=0.5, num_epochs=3, loss_type="mse")
    base.update(kw)
    return types.SimpleNamespace(**base)


def test_oracle_rays_and_pfm_match_reference():
    g = gold("g11_io")
    for tag, ndc in (("blender", False), ("llff", True)):
        H, W, focal = g[f"{tag}:hwf"].tolist()
        rays = IO.image_rays(g[f"{tag}:c2w"], int(H), int(W), focal, 2.0, 6.0, ndc)
        assert torch.equal(rays, g[f"{tag}:rays"])
    depth = g["pfm:depth"].numpy()
    assert IO.pfm_bytes(np.nan_to_num(depth)) == g["pfm:bytes"].numpy().tobytes()


def test_grid_points_order():
    p = IO.grid_points(3, (0, 1), (10, 11), (20, 21))
    assert p.shape == (27, 3)
    assert p[1].tolist() == [0.0, 10.0, 20.5] and p[3].tolist() == [0.5, 10.0, 20.0] and p[9].tolist() == [0.0, 10.5, 20.0]


@pytest.mark.gpu
@pytest.mark.parametrize("tag,ndc", [("blender", False), ("llff", True)])
def test_generate_rays_matches_reference(tag, ndc):
    import nerf_siren_b200 as nb
    g = gold("g11_io")
    H, W, focal = g[f"{tag}:hwf"].tolist()
    rays = nb.ops.generate_rays(g[f"{tag}:c2w"], int(H), int(W), focal, 2.0, 6.0, ndc).cpu()
    ref = g[f"{tag}:rays"]
    # fp32; the 3-term rotation and the norm are summed in a fixed fma order here, in MKL's order in the reference
    torch.testing.assert_close(rays, ref, rtol=2e-6, atol=2e-6)
    assert torch.equal(rays[:, 6:], ref[:, 6:])


@pytest.mark.gpu
def test_finish_image_quantise_flip_psnr():
    import nerf_siren_b200 as nb
    from nerf_siren_b200 import evaluation as ev
    g = gold("g11_io")
    H, W = 37, 53
    gen = torch.Generator().manual_seed(1)
    rgb = torch.rand(H * W, 3, generator=gen)
    rgb[0] = torch.tensor([0.0, 1.0, 0.999999])
    gt = torch.rand(H * W, 3, generator=gen)
    depth = g["pfm:depth"].reshape(-1).contiguous()
    u8, dp, se = nb.ops.finish_image(rgb.to(DEV), H, W, depth.to(DEV), gt.to(DEV))
    assert np.array_equal(u8.cpu().numpy(), IO.quantize(rgb.view(H, W, 3).numpy()))
    psnr = float(-10 * torch.log10(se / (3 * H * W)))
    assert abs(psnr - float(IO.psnr(rgb, gt))) < 1e-4
    import tempfile
    with tempfile.NamedTemporaryFile(suffix=".pfm") as f:
        ev.save_pfm(f.name, dp.cpu().numpy(), already_flipped=True)
        assert open(f.name, "rb").read() == g["pfm:bytes"].numpy().tobytes()          # byte-identical to the reference's file


@pytest.mark.gpu
def test_sigma_grid_matches_oracle():
    """extract_color_mesh.py:117-141 on the tcgen05 arm; bf16 tolerance 3e-2 of the sigma scale, same ordering as np.meshgrid."""
    import nerf_siren_b200 as nb
    from nerf_siren_b200 import evaluation as ev
    p = O.nerf_field(seed=21)
    m = nb.NeRF().to(DEV); m.load_state_dict(p)
    N, xr, yr, zr = 24, (-1.2, 1.2), (-1.0, 1.3), (-0.7, 1.1)
    got = ev.sigma_grid(m, nb.Embedding(3, 10), N, xr, yr, zr).cpu()
    pts = IO.grid_points(N, xr, yr, zr)
    ref = np.maximum(O.nerf_mlp(p, O.embed(pts, 10), True)[:, 0].numpy(), 0).reshape(N, N, N)
    assert got.shape == (N, N, N) and float(got.min()) >= 0
    scale = float(np.abs(ref).max())
    assert float(np.abs(got.numpy() - ref).max()) <= 3e-2 * scale


def test_oracle_vol_pack_known_answers():
    """extract_mesh.ipynb cell 7 by hand: voxel 1 (sigma <= 0) is dropped, rgb truncates, alpha = trunc(255 (1 - exp(-dx sigma)))."""
    v = np.array([[0.5, 0.25, 1.0, 100.0], [0.9, 0.9, 0.9, -3.0], [0.0, 0.999, 0.2, 1e4]], np.float32)
    got = IO.vol_pack(v, -1.2, 1.2, 512).reshape(-1, 2)
    assert got.dtype == np.uint32 and got[:, 0].tolist() == [0, 2]
    a0 = int(np.float32(255) * (np.float32(1) - np.exp(np.float32(-2.4 / 512) * np.float32(100.0))))
    assert got[0, 1] == (127 << 24) + (63 << 16) + (255 << 8) + a0
    assert got[1, 1] == (0 << 24) + (254 << 16) + (51 << 8) + 255


@pytest.mark.gpu
def test_vol_export_matches_oracle(tmp_path):
    """Row f4, `.vol` packing: the device compaction against the notebook's numpy on the SAME (N^3, 4) grid output -- index
    set and rgb bytes identical, alpha byte within 1 (numpy's float32 exp vs a correctly rounded one); and the grid's rgb /
    sigma against the fp32 field at bf16 tolerance."""
    import nerf_siren_b200 as nb
    from nerf_siren_b200 import _field, evaluation as ev, ops
    p = O.nerf_field(seed=22)
    m = nb.NeRF().to(DEV); m.load_state_dict(p)
    N, xr, yr, zr = 20, (-1.2, 1.2), (-1.2, 1.2), (-1.2, 1.2)
    emb = nb.Embedding(3, 10)
    rgbsigma = ops.sigma_grid(m._desc(False, 10, 4, True), _field.flat_params(m).detach(), N, xr, yr, zr, sigma_only=False)
    pts = IO.grid_points(N, xr, yr, zr)
    x = torch.cat([O.embed(pts, 10), O.embed(torch.zeros_like(pts), 4)], 1)
    ref = O.nerf_mlp(p, x, False).numpy()
    got = rgbsigma.cpu().numpy()
    def logit(v):
        v = np.clip(v.astype(np.float64), 1e-6, 1 - 1e-6)
        return np.log(v / (1 - v))
    lg, lr = logit(got[:, :3]), logit(ref[:, :3])                     # bf16 tolerance on the pre-sigmoid scale, as for sigma
    assert np.abs(lg - lr).max() <= 3e-2 * np.abs(lr).max() and np.abs(got[:, :3] - ref[:, :3]).mean() <= 1e-2
    assert np.abs(got[:, 3] - ref[:, 3]).max() <= 3e-2 * np.abs(ref[:, 3]).max()
    path = os.path.join(tmp_path, "scene.vol")
    pairs = ev.export_vol(path, m, emb, N, xr, yr, zr).cpu().numpy()
    want = IO.vol_pack(got, xr[0], xr[1], N).reshape(-1, 2)
    assert len(want) > 100 and len(want) < N ** 3                  # the case exercises both kept and dropped voxels
    assert pairs.shape == want.shape and np.array_equal(pairs[:, 0], want[:, 0])
    assert np.array_equal(pairs[:, 1] >> 8, want[:, 1] >> 8)
    assert np.abs((pairs[:, 1] & 255).astype(np.int64) - (want[:, 1] & 255).astype(np.int64)).max() <= 1
    assert open(path, "rb").read() == pairs.tobytes()
    # ragged sizes: a voxel count that is not a multiple of the 1024-voxel chunk, all-empty and empty inputs
    odd = rgbsigma[:1531].contiguous()
    assert np.array_equal(ops.pack_vol(odd, -2.4 / N).cpu().numpy()[:, 0], IO.vol_pack(got[:1531], xr[0], xr[1], N).reshape(-1, 2)[:, 0])
    none = rgbsigma.clone(); none[:, 3] = -1.0
    assert ops.pack_vol(none, -2.4 / N).shape == (0, 2)
    assert ops.pack_vol(rgbsigma[:0], -2.4 / N).shape == (0, 2)


@pytest.mark.gpu
def test_eg3d_renderer_glue():
    """Row f3: eg3d_training/eg3d_renderer.py:47-67 -- render() returns the six reference keys with the batch axis squeezed,
    sample() returns run_model's dict; both equal the ImportanceRenderer they wrap (same planes, decoder, random numbers)."""
    from nerf_siren_b200 import eg3d
    torch.manual_seed(5)
    kw = dict(eg3d.DEFAULT_RENDERING_KWARGS, depth_resolution=12, depth_resolution_importance=12, ray_start=0.5, ray_end=4.0, box_warp=3.0)
    r = eg3d.EG3D_Renderer(eg3d.LearnedPlanes(resolution=32), kw).to(DEV)
    n = 300
    o = torch.randn(n, 3, device=DEV) * 0.1 + torch.tensor([0.0, 0.0, -2.0], device=DEV)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=DEV) * 0.2 + torch.tensor([0.0, 0.0, 1.0], device=DEV), dim=-1)
    rc, u = torch.rand(1, n, 12, 1, device=DEV), torch.rand(1, n, 12, device=DEV)
    with torch.no_grad():
        out = r.render(-1, o, d, rand_coarse=rc, u=u)
        planes = r.plane_source().view(1, 3, 32, 32, 32)
        want = eg3d.ImportanceRenderer()(planes, r.decoder, o[None], d[None], kw, rand_coarse=rc, u=u)
        smp = r.sample(o[None] * 0.3, d[None])
    assert list(out) == ["rgb_coarse", "depth_coarse", "opacity_coarse", "rgb_fine", "depth_fine", "opacity_fine"]
    for k, w in zip(out, want):
        assert out[k].shape == w.shape[1:] and torch.equal(out[k], w[0])
    assert smp["rgb"].shape == (1, n, 3) and smp["sigma"].shape == (1, n, 1)
    assert "plane_source.planes" in dict(r.named_parameters())       # checkpoint prefix 'eg3d_renderer.' loads like system.py:25


def test_checkpoint_layout_and_schedulers(tmp_path):
    """Lightning key layout (utils/__init__.py:56-86 reads it back), MultiStepLR / cosine / poly as utils/__init__.py:33-52."""
    pytest.importorskip("nerf_siren_b200")
    from nerf_siren_b200 import system as S
    sd = {"state_dict": {"nerf_coarse.sigma.weight": torch.ones(1, 256), "nerf_fine.sigma.weight": torch.zeros(1, 256),
                         "nerf_fine.extra.weight": torch.zeros(1)}}
    path = os.path.join(tmp_path, "a.ckpt")
    torch.save(sd, path)
    got = S.extract_model_state_dict(path, "nerf_fine", prefixes_to_ignore=["extra"])
    assert list(got) == ["sigma.weight"] and float(got["sigma.weight"].sum()) == 0
    lin = torch.nn.Linear(2, 2)
    for name, want in (("steplr", [1.0, 1.0, 0.5]), ("poly", [1.0, (1 - 1 / 3) ** 0.9, (1 - 2 / 3) ** 0.9])):
        h = hp(lr=1.0, lr_scheduler=name)
        opt = S.get_optimizer(h, [lin]); sch = S.get_scheduler(h, opt)
        lrs = []
        for _ in range(3):
            lrs.append(S.get_learning_rate(opt)); opt.step(); sch.step()
        assert np.allclose(lrs, want)


@pytest.mark.gpu
def test_system_fit_trains_and_checkpoints(tmp_path):
    import nerf_siren_b200 as nb
    from nerf_siren_b200 import system as S
    torch.manual_seed(0)
    sys_ = S.NeRFSystem(hp(num_epochs=2), white_back=True).to(DEV)
    rays = O.blender_rays(512, seed=3).to(DEV)
    rgbs = torch.full((512, 3), 0.25, device=DEV)
    batches = [{"rays": rays[i:i + 256], "rgbs": rgbs[i:i + 256]} for i in (0, 256)] * 10
    hist = sys_.fit(batches, val_batches=[{"rays": rays[None], "rgbs": rgbs[None]}], ckpt_dir=str(tmp_path))
    assert len(hist) == 2 and hist[1]["val/loss"] < hist[0]["val/loss"] and hist[1]["val/psnr"] > hist[0]["val/psnr"]
    fresh = nb.NeRF().to(DEV)
    S.load_ckpt(fresh, os.path.join(tmp_path, "epoch=1.ckpt"), model_name="nerf_fine")
    for a, b in zip(fresh.parameters(), sys_.nerf_fine.parameters()):
        assert torch.equal(a, b)
    out = sys_(rays)                                     # system.py:199-223 result keys
    assert set(out) == {"rgb_coarse", "depth_coarse", "opacity_coarse", "rgb_fine", "depth_fine", "opacity_fine"}


@pytest.mark.gpu
def test_eg3d_system_fit_trains_and_checkpoints(tmp_path):
    """EG3DSystem (system.py:17-170 of the reference without Lightning): training_step through the fused render + the two
    backward kernels, loss_dict['mse'] on the six-key result, chunked validation, Lightning-layout checkpoint that
    ``hparams.pretrained`` reads back (system.py:24-25)."""
    from nerf_siren_b200 import eg3d, system as S
    from oracle import eg3d_oracle as E
    torch.manual_seed(0)
    kw = dict(eg3d.DEFAULT_RENDERING_KWARGS, depth_resolution=16, depth_resolution_importance=16, ray_start=2.2, ray_end=3.3, box_warp=1.0)
    sys_ = S.EG3DSystem(hp(num_epochs=2, lr=2e-2), plane_source=eg3d.LearnedPlanes(resolution=32), rendering_kwargs=kw).to(DEV)
    ro, rd = E.camera_rays(9000, seed=2)
    rays = torch.cat([ro[0], rd[0], torch.zeros(9000, 2)], 1).to(DEV)
    rgbs = (torch.rand(9000, 3, generator=torch.Generator().manual_seed(1)) * 0.4 + 0.3).to(DEV)
    batches = [{"rays": rays[i:i + 1024], "rgbs": rgbs[i:i + 1024]} for i in range(0, 8192, 1024)]
    hist = sys_.fit(batches, val_batches=[{"rays": rays[None], "rgbs": rgbs[None]}], ckpt_dir=str(tmp_path))   # 9000 rays: 3 chunks
    assert len(hist) == 2 and hist[1]["val/loss"] < hist[0]["val/loss"] and hist[1]["val/psnr"] > hist[0]["val/psnr"]
    assert hist[1]["train/loss"] < hist[0]["train/loss"]
    out = sys_(rays[:100])
    assert set(out) == {"rgb_coarse", "depth_coarse", "opacity_coarse", "rgb_fine", "depth_fine", "opacity_fine"}
    assert out["rgb_fine"].shape == (100, 3) and out["rgb_fine"].requires_grad
    again = S.EG3DSystem(hp(pretrained=os.path.join(tmp_path, "epoch=1.ckpt")), plane_source=eg3d.LearnedPlanes(resolution=32),
                         rendering_kwargs=kw).to(DEV)
    for (ka, a), (kb, b) in zip(again.eg3d_renderer.state_dict().items(), sys_.eg3d_renderer.state_dict().items()):
        assert ka == kb and torch.equal(a, b), ka
