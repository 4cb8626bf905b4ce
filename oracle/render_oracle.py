"""CPU oracle for the render_rays hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A torch-CPU / fp32 restatement of the reference algorithm (Freedomcls/nerf-siren), written
functionally: every function takes plain tensors and a ``{state_dict name: tensor}`` mapping
instead of nn.Modules, and every random tensor the reference draws internally (``torch.rand`` /
``torch.randn``) is an explicit argument so the CUDA path can be fed the very same numbers.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module, and only as the checker or the timed CPU baseline.  The product
package (``nerf_siren_b200``) never imports it and has no CPU fallback.

Parity pin: ``oracle/make_goldens.py`` imports the real reference from ``/root/reference`` (in the
build container only), runs both on the same seeded inputs, asserts bit-equality (forward) and
writes the reference's outputs to ``tests/golden/*.npz``.  ``tests/test_oracle_golden.py``
re-checks this module against those files wherever the tests run.  The reference itself ships no
golden vectors for this path -- the committed fixtures ARE reference outputs.

Citations are ``file:line`` into the reference tree.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

Params = Dict[str, torch.Tensor]


# ----------------------------------------------------------------------------------------------
# a1  positional encoding                                    models/nerf.py:5-19 (bands), :33-38
# ----------------------------------------------------------------------------------------------
def freq_bands(n_freqs: int, logscale: bool = True) -> torch.Tensor:
    """models/nerf.py:16-19 -- 2**linspace(0, N-1, N) (exact powers of two) or a linear ramp."""
    if n_freqs == 0:
        return torch.zeros(0)
    if logscale:
        return 2 ** torch.linspace(0, n_freqs - 1, n_freqs)
    return torch.linspace(1, 2 ** (n_freqs - 1), n_freqs)


def embed(x: torch.Tensor, n_freqs: int, logscale: bool = True) -> torch.Tensor:
    """[x, sin(f0 x), cos(f0 x), sin(f1 x), ...]; channel order of models/nerf.py:33-38."""
    pieces = [x]
    for f in freq_bands(n_freqs, logscale):
        fx = f * x
        pieces.append(torch.sin(fx))
        pieces.append(torch.cos(fx))
    return torch.cat(pieces, dim=-1)


# ----------------------------------------------------------------------------------------------
# a2 / a3  NeRF and NeRF_3D MLPs            models/nerf.py:83-124, models/nerf_cls.py:70-112
# ----------------------------------------------------------------------------------------------
def _lin(p: Params, name: str, x: torch.Tensor) -> torch.Tensor:
    return F.linear(x, p[name + ".weight"], p[name + ".bias"])


def nerf_mlp(p: Params, x: torch.Tensor, sigma_only: bool = False, D: int = 8,
             in_xyz: int = 63, in_dir: int = 27, skips: Sequence[int] = (4,)) -> torch.Tensor:
    """NeRF.forward (models/nerf.py:100-124) / NeRF_3D.forward (models/nerf_cls.py:87-112).

    The class head is applied iff the mapping holds ``parse.0.weight`` (NeRF_3D).  Output column
    order is [r, g, b, sigma(, cls...)]; sigma is the raw pre-activation (nerf.py:112,122).
    """
    xyz_in = x[..., :in_xyz]
    h = xyz_in
    for i in range(D):
        if i in skips:                                   # nerf.py:108-109: cat BEFORE layer i+1
            h = torch.cat([xyz_in, h], dim=-1)
        h = torch.relu(_lin(p, f"xyz_encoding_{i + 1}.0", h))
    sigma = _lin(p, "sigma", h)                          # reads post-ReLU layer-D output
    if sigma_only:
        return sigma
    feat = _lin(p, "xyz_encoding_final", h)              # no activation (nerf.py:116)
    dir_in = x[..., in_xyz:in_xyz + in_dir]
    hd = torch.relu(_lin(p, "dir_encoding.0", torch.cat([feat, dir_in], dim=-1)))
    rgb = torch.sigmoid(_lin(p, "rgb.0", hd))
    cols = [rgb, sigma]
    if "parse.0.weight" in p:                            # nerf_cls.py:56-61,104: Linear,Linear,Sigmoid
        cols.append(torch.sigmoid(_lin(p, "parse.1", _lin(p, "parse.0", feat))))
    return torch.cat(cols, dim=-1)


# ----------------------------------------------------------------------------------------------
# a4  SIREN / FiLM MLP                                                models/nerf.py:142-216
# ----------------------------------------------------------------------------------------------
def siren_mlp(p: Params, pts: torch.Tensor, frequencies: torch.Tensor, phase_shifts: torch.Tensor,
              ray_directions: torch.Tensor, hidden: int = 256, box: float = 51.0) -> torch.Tensor:
    """SemanticNeRF.forward_with_frequencies_phase_shifts (models/nerf.py:201-216).

    pts, ray_directions: (Bb, P, 3); frequencies, phase_shifts: (Bb, 9*hidden).
    """
    frequencies = frequencies * 15 + 30                                   # :202
    h = pts * (2.0 / box)                                                  # UniformBoxWarp :134-140
    n_film = 0
    while f"network.{n_film}.layer.weight" in p:
        n_film += 1
    for l in range(n_film):
        a, b = l * hidden, (l + 1) * hidden
        z = _lin(p, f"network.{l}.layer", h)
        h = torch.sin(frequencies[..., a:b].unsqueeze(1) * z + phase_shifts[..., a:b].unsqueeze(1))
    sigma = _lin(p, "final_layer", h)
    zc = _lin(p, "color_layer_sine.layer", torch.cat([ray_directions, h], dim=-1))      # :213
    hc = torch.sin(frequencies[..., -hidden:].unsqueeze(1) * zc + phase_shifts[..., -hidden:].unsqueeze(1))
    rgb = torch.sigmoid(_lin(p, "color_layer_linear.0", hc))
    return torch.cat([rgb, sigma], dim=-1)


def siren_init(hidden: int = 256, seed: int = 0) -> Params:
    """Parameter set with the reference's init (models/nerf.py:126-132,153-157,187-191):
    U(+-sqrt(6/in)/25) everywhere, first FiLM layer U(+-1/in); biases keep nn.Linear's default."""
    g = torch.Generator().manual_seed(seed)

    def lin(out_f, in_f, first=False):
        bound = (1.0 / in_f) if first else math.sqrt(6.0 / in_f) / 25.0
        w = (torch.rand(out_f, in_f, generator=g) * 2 - 1) * bound
        bb = 1.0 / math.sqrt(in_f)
        b = (torch.rand(out_f, generator=g) * 2 - 1) * bb
        return w, b

    p: Params = {}
    for l in range(8):
        w, b = lin(hidden, 3 if l == 0 else hidden, first=(l == 0))
        p[f"network.{l}.layer.weight"], p[f"network.{l}.layer.bias"] = w, b
    p["final_layer.weight"], p["final_layer.bias"] = lin(1, hidden)
    p["color_layer_sine.layer.weight"], p["color_layer_sine.layer.bias"] = lin(hidden, hidden + 3)
    p["color_layer_linear.0.weight"], p["color_layer_linear.0.bias"] = lin(3, hidden)
    return p


# ----------------------------------------------------------------------------------------------
# a5  stratified depth samples                                   models/rendering.py:198-225
# ----------------------------------------------------------------------------------------------
def stratified_z(rays: torch.Tensor, n_samples: int, use_disp: bool = False, perturb: float = 0.0,
                 perturb_rand: Optional[torch.Tensor] = None) -> torch.Tensor:
    """z_vals (N, n_samples).  ``perturb_rand`` is the U[0,1) tensor of rendering.py:221 (before
    the multiplication by ``perturb``); it is required iff perturb > 0."""
    near, far = rays[:, 6:7], rays[:, 7:8]
    t = torch.linspace(0, 1, n_samples)
    if not use_disp:
        z = near * (1 - t) + far * t
    else:
        z = 1 / (1 / near * (1 - t) + 1 / far * t)
    z = z.expand(rays.shape[0], n_samples)
    if perturb > 0:
        mid = 0.5 * (z[:, :-1] + z[:, 1:])
        upper = torch.cat([mid, z[:, -1:]], dim=-1)
        lower = torch.cat([z[:, :1], mid], dim=-1)
        z = lower + (upper - lower) * (perturb * perturb_rand)
    return z


# ----------------------------------------------------------------------------------------------
# a7 / a10  alpha compositing                                    models/rendering.py:161-190,379
# ----------------------------------------------------------------------------------------------
def composite(sigmas: torch.Tensor, z: torch.Tensor, rays_d: torch.Tensor, noise: Optional[torch.Tensor],
              noise_std: float, rgbs: Optional[torch.Tensor] = None, white_back: bool = False,
              extra: Optional[torch.Tensor] = None):
    """Returns ``weights`` if rgbs is None, else (rgb, depth, weights[, extra_map]).

    sigmas, z: (N,S); rgbs: (N,S,3); extra: (N,S,C) per-sample class scores composited as
    sum_i w_i * extra_i (rendering.py:379).  ``noise`` is the N(0,1) draw of rendering.py:170.
    """
    deltas = z[:, 1:] - z[:, :-1]
    deltas = torch.cat([deltas, 1e10 * torch.ones_like(deltas[:, :1])], dim=-1)      # :163-164
    deltas = deltas * torch.norm(rays_d.unsqueeze(1), dim=-1)                         # :168
    nz = noise * noise_std if noise is not None else torch.zeros_like(sigmas)
    alphas = 1 - torch.exp(-deltas * torch.relu(sigmas + nz))                         # :173
    shifted = torch.cat([torch.ones_like(alphas[:, :1]), 1 - alphas + 1e-10], dim=-1)
    weights = alphas * torch.cumprod(shifted, dim=-1)[:, :-1]                         # :176-177
    if rgbs is None:
        return weights
    wsum = weights.sum(1)
    rgb = torch.sum(weights.unsqueeze(-1) * rgbs, dim=-2)
    depth = torch.sum(weights * z, dim=-1)
    if white_back:
        rgb = rgb + 1 - wsum.unsqueeze(-1)                                            # :187-188
    if extra is None:
        return rgb, depth, weights
    return rgb, depth, weights, torch.sum(weights.unsqueeze(-1) * extra, dim=-2)


# ----------------------------------------------------------------------------------------------
# a8  hierarchical resampling                                      models/rendering.py:22-67
# ----------------------------------------------------------------------------------------------
def linspace01(n: int) -> torch.Tensor:
    """torch.linspace(0, 1, n) in fp32 -- the deterministic ``u`` (rendering.py:44) and z_steps."""
    return torch.linspace(0, 1, n)


def sample_pdf(bins: torch.Tensor, weights: torch.Tensor, n_importance: int, det: bool = False,
               eps: float = 1e-5, u: Optional[torch.Tensor] = None, return_aux: bool = False):
    """Inverse-CDF sampling.  ``u`` (N, n_importance) replaces the torch.rand draw of :47 when
    det is False.  With return_aux also returns (cdf, u, inds) for the index-exactness tests."""
    n_rays, n_w = weights.shape
    w = weights + eps
    pdf = w / torch.sum(w, dim=-1, keepdim=True)
    cdf = torch.cumsum(pdf, dim=-1)
    cdf = torch.cat([torch.zeros_like(cdf[:, :1]), cdf], dim=-1)
    if det:
        u = linspace01(n_importance).expand(n_rays, n_importance)
    elif u is None:
        u = torch.rand(n_rays, n_importance)
    u = u.contiguous()
    inds = torch.searchsorted(cdf.detach(), u, right=True)                            # :54
    below = torch.clamp_min(inds - 1, 0)
    above = torch.clamp_max(inds, n_w)
    c0, c1 = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    b0, b1 = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = c1 - c0
    denom = torch.where(denom < eps, torch.ones_like(denom), denom)                   # :63
    samples = b0 + (u - c0) / denom * (b1 - b0)
    if return_aux:
        return samples, cdf, u, inds
    return samples


def searchsorted_rows(a: np.ndarray, v: np.ndarray, side: str = "left") -> np.ndarray:
    """Row-batched numpy.searchsorted with 1-row broadcast -- the contract
    torchsearchsorted/test/test_searchsorted.py:34-44 pins (numpy oracle: .../utils.py:4-15)."""
    rows = max(a.shape[0], v.shape[0])
    out = np.empty((rows, v.shape[1]), dtype=np.int64)
    for r in range(rows):
        out[r] = np.searchsorted(a[r if a.shape[0] > 1 else 0], v[r if v.shape[0] > 1 else 0], side=side)
    return out


# ----------------------------------------------------------------------------------------------
# a6 / a9  render_rays orchestration                               models/rendering.py:70-262
# ----------------------------------------------------------------------------------------------
def _run_field(field, xyz: torch.Tensor, rays_d: torch.Tensor, dir_emb: torch.Tensor, n_freq_xyz: int,
               sigma_only: bool, chunk: int) -> torch.Tensor:
    """The point-chunk loop of rendering.py:129-152.  ``field(x_embedded, sigma_only)`` -> (B,C)."""
    n_rays, s, _ = xyz.shape
    flat = xyz.reshape(-1, 3)
    rep = None if sigma_only else torch.repeat_interleave(dir_emb, repeats=s, dim=0)
    outs = []
    for i in range(0, flat.shape[0], chunk):
        e = embed(flat[i:i + chunk], n_freq_xyz)
        if not sigma_only:
            e = torch.cat([e, rep[i:i + chunk]], dim=1)
        outs.append(field(e, sigma_only))
    return torch.cat(outs, dim=0).view(n_rays, s, -1)


def render_rays(fields, rays: torch.Tensor, n_samples: int = 64, use_disp: bool = False,
                perturb: float = 0.0, noise_std: float = 1.0, n_importance: int = 0,
                chunk: int = 1024 * 32, white_back: bool = False, test_time: bool = False,
                n_freq_xyz: int = 10, n_freq_dir: int = 4,
                rng: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
    """models/rendering.py:70-262 with explicit randomness.

    fields : [coarse, fine] callables ``f(x_embedded, sigma_only) -> (B, 1 | 4 | 4+CLS)``.
    rng    : {'perturb_rand': U(N,Sc), 'noise_coarse': N(N,Sc), 'u': U(N,Si), 'noise_fine': N(N,Sc+Si)}
             -- the four draws of one call, in the reference's order.  Entries that the
             configuration does not consume may be absent.  Extra output channels (>4) are
             composited into ``cls_coarse`` / ``cls_fine`` (rendering.py:379,447,469).
    """
    rng = rng or {}
    rays_o, rays_d = rays[:, 0:3], rays[:, 3:6]
    dir_emb = embed(rays_d, n_freq_dir)
    z = stratified_z(rays, n_samples, use_disp, perturb, rng.get("perturb_rand"))
    out: Dict[str, torch.Tensor] = {}

    def shade(field, z_vals, noise, sigma_only):
        xyz = rays_o.unsqueeze(1) + rays_d.unsqueeze(1) * z_vals.unsqueeze(2)
        raw = _run_field(field, xyz, rays_d, dir_emb, n_freq_xyz, sigma_only, chunk)
        if sigma_only:
            return (composite(raw[..., 0], z_vals, rays_d, noise, noise_std),)
        extra = raw[..., 4:] if raw.shape[-1] > 4 else None
        return composite(raw[..., 3], z_vals, rays_d, noise, noise_std, raw[..., :3], white_back, extra)

    res = shade(fields[0], z, rng.get("noise_coarse"), test_time)
    w_coarse = res[0] if test_time else res[2]
    if not test_time:
        out["rgb_coarse"], out["depth_coarse"] = res[0], res[1]
        if len(res) == 4:
            out["cls_coarse"] = res[3]
    out["opacity_coarse"] = w_coarse.sum(1)

    if n_importance > 0:
        mid = 0.5 * (z[:, :-1] + z[:, 1:])
        z_new = sample_pdf(mid, w_coarse[:, 1:-1], n_importance, det=(perturb == 0),
                           u=rng.get("u")).detach()
        z_all, _ = torch.sort(torch.cat([z, z_new], dim=-1), dim=-1)
        res = shade(fields[1], z_all, rng.get("noise_fine"), False)
        out["rgb_fine"], out["depth_fine"], out["opacity_fine"] = res[0], res[1], res[2].sum(1)
        if len(res) == 4:
            out["cls_fine"] = res[3]
        out["_z_fine"] = z_all                      # oracle-only extras (underscore-prefixed)
        out["_weights_fine"] = res[2]
    out["_z_coarse"] = z
    out["_weights_coarse"] = w_coarse
    return out


def mse_loss(out: Dict[str, torch.Tensor], target: torch.Tensor) -> torch.Tensor:
    """losses.py:15-20 -- mean squared error of rgb_coarse (+ rgb_fine when present)."""
    loss = torch.mean((out["rgb_coarse"] - target) ** 2)
    if "rgb_fine" in out:
        loss = loss + torch.mean((out["rgb_fine"] - target) ** 2)
    return loss


# ----------------------------------------------------------------------------------------------
# parameter helpers shared by tests / bench (default nn.Linear init, reference key names)
# ----------------------------------------------------------------------------------------------
def nerf_param_shapes(D: int = 8, W: int = 256, in_xyz: int = 63, in_dir: int = 27,
                      skips: Sequence[int] = (4,), n_cls: int = 0):
    """(name, shape) in the reference's state_dict order (models/nerf.py:59-81,
    models/nerf_cls.py:32-61)."""
    shapes = []
    for i in range(D):
        fan_in = in_xyz if i == 0 else (W + in_xyz if i in skips else W)
        shapes += [(f"xyz_encoding_{i + 1}.0.weight", (W, fan_in)), (f"xyz_encoding_{i + 1}.0.bias", (W,))]
    shapes += [("xyz_encoding_final.weight", (W, W)), ("xyz_encoding_final.bias", (W,))]
    shapes += [("dir_encoding.0.weight", (W // 2, W + in_dir)), ("dir_encoding.0.bias", (W // 2,))]
    shapes += [("sigma.weight", (1, W)), ("sigma.bias", (1,))]
    shapes += [("rgb.0.weight", (3, W // 2)), ("rgb.0.bias", (3,))]
    if n_cls:
        shapes += [("parse.0.weight", (2 * W, W)), ("parse.0.bias", (2 * W,)),
                   ("parse.1.weight", (n_cls, 2 * W)), ("parse.1.bias", (n_cls,))]
    return shapes


def nerf_init(seed: int = 0, gain: float = 1.0, **kw) -> Params:
    """Default nn.Linear-style init (kaiming_uniform a=sqrt(5) == U(+-1/sqrt(fan_in)) for weight and
    bias) from a private generator; ``gain`` scales the weights to get a non-flat field."""
    g = torch.Generator().manual_seed(seed)
    p: Params = {}
    for name, shape in nerf_param_shapes(**kw):
        fan_in = shape[1] if len(shape) == 2 else None
        if fan_in is None:                      # bias follows its weight: same fan_in bound
            fan_in = p[name.replace(".bias", ".weight")].shape[1]
        bound = 1.0 / math.sqrt(fan_in)
        t = (torch.rand(*shape, generator=g) * 2 - 1) * bound
        if len(shape) == 2:
            t = t * gain
        p[name] = t
    return p


def nerf_field(seed: int = 0, gain: float = 3.0, sigma_mean: float = 0.5, sigma_std: float = 3.0, **kw) -> Params:
    """A non-degenerate synthetic field for parity tests: ``nerf_init`` weights whose sigma head is affinely
    re-calibrated so that raw sigma has the given mean / std over a probe set of Blender-shaped samples (a plain
    random init yields sigma < 0 almost everywhere, i.e. an empty scene that would hide errors)."""
    p = nerf_init(seed=seed, gain=gain, **kw)
    rays = blender_rays(32, seed=99)
    z = stratified_z(rays, 16)
    xyz = (rays[:, None, :3] + rays[:, None, 3:6] * z[..., None]).reshape(-1, 3)
    with torch.no_grad():
        s = nerf_mlp({k: v for k, v in p.items() if not k.startswith("parse")}, embed(xyz, 10), sigma_only=True)
        k = sigma_std / float(s.std())
        p["sigma.weight"] = p["sigma.weight"] * k
        p["sigma.bias"] = (p["sigma.bias"] - float(s.mean())) * k + sigma_mean
    return p


def blender_rays(n: int, seed: int = 0, hw: int = 800, radius: float = 4.0311, near: float = 2.0,
                 far: float = 6.0) -> torch.Tensor:
    """Synthetic Blender-shaped rays (datasets/blender.py:30-42,64-69; datasets/ray_utils.py:17-50):
    pinhole camera on a sphere looking at the origin, unit directions, near 2 / far 6."""
    g = torch.Generator().manual_seed(seed)
    focal = 0.5 * hw / math.tan(0.5 * 0.6911112070083618)
    th = torch.rand(1, generator=g).item() * 2 * math.pi
    ph = (0.15 + 0.5 * torch.rand(1, generator=g).item()) * math.pi / 2
    cam = torch.tensor([math.cos(th) * math.cos(ph), math.sin(th) * math.cos(ph), math.sin(ph)]) * radius
    fwd = -cam / cam.norm()
    right = torch.linalg.cross(fwd, torch.tensor([0.0, 0.0, 1.0]))
    right = right / right.norm()
    up = torch.linalg.cross(right, fwd)
    c2w = torch.stack([right, up, -fwd], dim=1)                      # columns: x, y, z(back)
    ij = torch.randint(0, hw * hw, (n,), generator=g)
    i, j = (ij % hw).float(), (ij // hw).float()
    d_cam = torch.stack([(i - hw / 2) / focal, -(j - hw / 2) / focal, -torch.ones(n)], dim=-1)
    d = d_cam @ c2w.T
    d = d / d.norm(dim=-1, keepdim=True)
    o = cam.expand(n, 3)
    return torch.cat([o, d, near * torch.ones(n, 1), far * torch.ones(n, 1)], dim=1).contiguous()


def llff_rays(n: int, seed: int = 0, w: int = 504, h: int = 378, focal: float = 407.0) -> torch.Tensor:
    """Synthetic LLFF/NDC-shaped rays (datasets/ray_utils.py:53-94, datasets/llff.py:235-250):
    origins on the z=-1 plane, un-normalised directions, near 0 / far 1."""
    g = torch.Generator().manual_seed(seed)
    ij = torch.randint(0, w * h, (n,), generator=g)
    i, j = (ij % w).float(), (ij // w).float()
    d = torch.stack([(i - w / 2) / focal, -(j - h / 2) / focal, -torch.ones(n)], dim=-1)
    o = 0.05 * torch.randn(3, generator=g).expand(n, 3)
    nearp = 1.0
    t = -(nearp + o[:, 2]) / d[:, 2]
    o = o + t[:, None] * d
    ox, oy, oz = o[:, 0] / o[:, 2], o[:, 1] / o[:, 2], o[:, 2]
    o_ndc = torch.stack([-focal / (w / 2) * ox, -focal / (h / 2) * oy, 1 + 2 * nearp / oz], dim=-1)
    d_ndc = torch.stack([-focal / (w / 2) * (d[:, 0] / d[:, 2] - ox), -focal / (h / 2) * (d[:, 1] / d[:, 2] - oy),
                         1 - o_ndc[:, 2]], dim=-1)
    return torch.cat([o_ndc, d_ndc, torch.zeros(n, 1), torch.ones(n, 1)], dim=1).contiguous()


# ----------------------------------------------------------------------------------------------
# bf16 tensor-core arithmetic, emulated  (checker for the tcgen05 arm's OWN arithmetic)
# ----------------------------------------------------------------------------------------------
def _r16(t: torch.Tensor) -> torch.Tensor:
    """Round to bf16 (nearest even) and back: the value a bf16 UMMA operand carries."""
    return t.to(torch.bfloat16).to(torch.float32)


def nerf_mlp_bf16(p: Params, x: torch.Tensor, d_out: Optional[torch.Tensor] = None, D: int = 8, in_xyz: int = 63,
                  in_dir: int = 27, skips: Sequence[int] = (4,)):
    """models/nerf.py:100-124 evaluated the way the tensor-core kernels evaluate it: bf16 weights and bf16
    layer inputs, fp32 accumulation, fp32 bias / ReLU / heads on the un-rounded accumulators; and, when
    ``d_out`` (B,4) is given, the manual backward with bf16 dZ and gates taken from the bf16 activations.
    Returns ``out`` or ``(out, {param name: grad})``.  The fp32 reference differs from this by bf16 rounding
    (and, for gradients, by ReLU gates that flip under that rounding); the CUDA kernels must match THIS to
    fp32-summation noise, which separates "bf16 is coarse" from "the kernel is wrong"."""
    W = lambda n: _r16(p[n + ".weight"])
    b = lambda n: p[n + ".bias"]
    X = _r16(x[:, :in_xyz])
    Xd = _r16(x[:, in_xyz:in_xyz + in_dir])
    acts, h, h32 = [], X, None                       # acts[i] = bf16 input of trunk layer i
    for i in range(D):
        inp = torch.cat([X, h], -1) if i in skips else h
        acts.append(inp)
        h32 = torch.relu(inp @ W(f"xyz_encoding_{i + 1}.0").T + b(f"xyz_encoding_{i + 1}.0"))
        h = _r16(h32)
    hD = h
    sigma = h32 @ p["sigma.weight"].T + p["sigma.bias"]
    feat = _r16(hD @ W("xyz_encoding_final").T + b("xyz_encoding_final"))
    din = torch.cat([feat, Xd], -1)
    hd32 = torch.relu(din @ W("dir_encoding.0").T + b("dir_encoding.0"))
    hd = _r16(hd32)
    rgb = torch.sigmoid(hd32 @ p["rgb.0.weight"].T + p["rgb.0.bias"])
    out = torch.cat([rgb, sigma], -1)
    if d_out is None:
        return out
    g = {}
    gr = d_out[:, :3] * rgb * (1 - rgb)
    dsig = d_out[:, 3:4]
    g["rgb.0.weight"], g["rgb.0.bias"] = gr.T @ hd, gr.sum(0)
    g["sigma.weight"], g["sigma.bias"] = dsig.T @ hD, dsig.sum(0)
    dz = _r16((gr @ p["rgb.0.weight"]) * (hd != 0))
    g["dir_encoding.0.weight"], g["dir_encoding.0.bias"] = dz.T @ din, dz.sum(0)
    dz = _r16(dz @ W("dir_encoding.0")[:, :feat.shape[1]])
    g["xyz_encoding_final.weight"], g["xyz_encoding_final.bias"] = dz.T @ hD, dz.sum(0)
    dh = dz @ W("xyz_encoding_final") + dsig @ p["sigma.weight"]
    for i in range(D - 1, -1, -1):
        name = f"xyz_encoding_{i + 1}.0"
        out_act = hD if i == D - 1 else (acts[i + 1][:, in_xyz:] if (i + 1) in skips else acts[i + 1])
        dz = _r16(dh * (out_act != 0))
        g[name + ".weight"], g[name + ".bias"] = dz.T @ acts[i], dz.sum(0)
        if i > 0:
            Wi = W(name)
            dh = dz @ (Wi[:, in_xyz:] if i in skips else Wi)
    return out, g
