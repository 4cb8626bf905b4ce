"""Drop-in for the reference's ``models/nerf.py``: same classes, constructor arguments, attributes and
state_dict keys, so checkpoints and ``utils.load_ckpt`` prefixes keep working -- but
``forward`` runs the fused sm_100a kernels of libnerf_b200.so instead of torch ops.

Reference: models/nerf.py:4-38 (Embedding), :41-124 (NeRF), :126-157 (SIREN inits, FiLMLayer),
:159-216 (SemanticNeRF).  CUDA tensors only; there is no CPU path.
"""
import math

import numpy as np  # the reference forgets this import (models/nerf.py:131); SemanticNeRF needs it
import torch
from torch import nn

from .. import _field, _lib, ops
from .._lib import MlpDesc

__all__ = ["Embedding", "NeRF", "frequency_init", "UniformBoxWarp", "FiLMLayer", "first_layer_film_sine_init",
           "SemanticNeRF", "SirenField"]


class _EmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, n_freqs, logscale, bands):
        ctx.save_for_backward(x, bands)
        return ops.embed(x, n_freqs, logscale)

    @staticmethod
    def backward(ctx, g):          # rarely needed (pose refinement); plain tensor algebra on the device
        x, bands = ctx.saved_tensors
        c = x.shape[1]
        gx = g[:, :c].clone()
        for k, f in enumerate(bands.tolist()):
            gs, gc = g[:, c + 2 * k * c: c + (2 * k + 1) * c], g[:, c + (2 * k + 1) * c: c + (2 * k + 2) * c]
            gx += f * (gs * torch.cos(f * x) - gc * torch.sin(f * x))
        return gx, None, None, None


class Embedding(nn.Module):
    """x -> (x, sin(2^k x), cos(2^k x), ...), models/nerf.py:4-38."""

    def __init__(self, in_channels, N_freqs, logscale=True):
        super().__init__()
        self.N_freqs = N_freqs
        self.in_channels = in_channels
        self.funcs = [torch.sin, torch.cos]
        self.out_channels = in_channels * (len(self.funcs) * N_freqs + 1)
        self.logscale = bool(logscale)
        if logscale:
            self.freq_bands = 2 ** torch.linspace(0, N_freqs - 1, N_freqs)
        else:
            self.freq_bands = torch.linspace(1, 2 ** (N_freqs - 1), N_freqs)

    def forward(self, x):
        if x.dim() != 2 or x.shape[1] != self.in_channels:
            raise ValueError(f"Embedding expects (B, {self.in_channels}), got {tuple(x.shape)}")
        if not x.is_cuda:
            raise RuntimeError("Embedding input must be a CUDA tensor (nerf_siren_b200 has no CPU path)")
        x = x.contiguous().float()
        if x.requires_grad:
            return _EmbedFn.apply(x, self.N_freqs, self.logscale, self.freq_bands)
        return ops.embed(x, self.N_freqs, self.logscale)


class NeRF(nn.Module):
    """models/nerf.py:41-124.  Parameters are registered through the same nn.Linear containers, in the same
    order, as the reference (identical names, shapes, default init and RNG consumption)."""

    CLS = 0

    def __init__(self, D=8, W=256, in_channels_xyz=63, in_channels_dir=27, skips=[4]):
        super().__init__()
        self.D, self.W = D, W
        self.in_channels_xyz, self.in_channels_dir = in_channels_xyz, in_channels_dir
        self.skips = skips
        for i in range(D):
            fan_in = in_channels_xyz if i == 0 else (W + in_channels_xyz if i in skips else W)
            setattr(self, f"xyz_encoding_{i + 1}", nn.Sequential(nn.Linear(fan_in, W), nn.ReLU(True)))
        self.xyz_encoding_final = nn.Linear(W, W)
        self.dir_encoding = nn.Sequential(nn.Linear(W + in_channels_dir, W // 2), nn.ReLU(True))
        self.sigma = nn.Linear(W, 1)
        self.rgb = nn.Sequential(nn.Linear(W // 2, 3), nn.Sigmoid())

    # -- description handed to the C ABI ---------------------------------------------------------------
    def _desc(self, embedded: bool, n_freq_xyz: int = 10, n_freq_dir: int = 4, logscale: bool = True) -> MlpDesc:
        mask = 0
        for s in self.skips:
            if 0 < s < self.D:
                mask |= 1 << s
        d = MlpDesc(kind=_lib.KIND_NERF, D=self.D, W=self.W, n_freq_xyz=n_freq_xyz, n_freq_dir=n_freq_dir,
                    skip_mask=mask, n_cls=self.CLS, logscale=int(logscale), box=0.0, in_xyz=0, in_dir=0)
        if embedded:
            d.in_xyz, d.in_dir = self.in_channels_xyz, self.in_channels_dir
        return d

    def _film(self):
        return None

    def forward(self, x, sigma_only=False):
        """x: (B, in_channels_xyz [+ in_channels_dir]) embedded input -> (B,1) sigma or (B,4[+CLS])."""
        want = self.in_channels_xyz + (0 if sigma_only else self.in_channels_dir)
        if x.dim() != 2 or x.shape[1] != want:
            raise ValueError(f"expected input of shape (B, {want}), got {tuple(x.shape)}")
        return _field.run_embedded(self, self._desc(True), x, bool(sigma_only))


# ------------------------------------------------------------------------------------------------------
# SIREN / FiLM (models/nerf.py:126-216)
# ------------------------------------------------------------------------------------------------------
def frequency_init(freq):
    def init(m):
        with torch.no_grad():
            if isinstance(m, nn.Linear):
                num_input = m.weight.size(-1)
                m.weight.uniform_(-np.sqrt(6 / num_input) / freq, np.sqrt(6 / num_input) / freq)
    return init


def first_layer_film_sine_init(m):
    with torch.no_grad():
        if isinstance(m, nn.Linear):
            num_input = m.weight.size(-1)
            m.weight.uniform_(-1 / num_input, 1 / num_input)


class UniformBoxWarp(nn.Module):
    def __init__(self, sidelength):
        super().__init__()
        self.scale_factor = 2 / sidelength

    def forward(self, coordinates):
        return coordinates * self.scale_factor


class FiLMLayer(nn.Module):
    """Parameter holder for one sin(freq * (Wx+b) + phase) layer (models/nerf.py:142-151).  Calling it on its
    own is supported for small inputs through plain tensor algebra on the device; the fused path never does."""

    def __init__(self, input_dim, hidden_dim):
        super().__init__()
        self.layer = nn.Linear(input_dim, hidden_dim)

    def forward(self, x, freq, phase_shift):
        x = self.layer(x)
        return torch.sin(freq.unsqueeze(1).expand_as(x) * x + phase_shift.unsqueeze(1).expand_as(x))


class SemanticNeRF(nn.Module):
    """models/nerf.py:159-216 (pi-GAN TALLSIREN + UniformBoxWarp(51)); same parameter names and inits."""

    def __init__(self, input_dim=2, z_dim=100, hidden_dim=256, output_dim=1, device=None):
        super().__init__()
        self.device = device
        self.input_dim, self.z_dim, self.hidden_dim, self.output_dim = input_dim, z_dim, hidden_dim, output_dim
        self.network = nn.ModuleList([FiLMLayer(3 if i == 0 else hidden_dim, hidden_dim) for i in range(8)])
        self.final_layer = nn.Linear(hidden_dim, 1)
        self.color_layer_sine = FiLMLayer(hidden_dim + 3, hidden_dim)
        self.color_layer_linear = nn.Sequential(nn.Linear(hidden_dim, 3))
        self.network.apply(frequency_init(25))
        self.final_layer.apply(frequency_init(25))
        self.color_layer_sine.apply(frequency_init(25))
        self.color_layer_linear.apply(frequency_init(25))
        self.network[0].apply(first_layer_film_sine_init)
        self.gridwarper = UniformBoxWarp(51)

    def _desc(self, embedded: bool = True, *a, **k) -> MlpDesc:
        return MlpDesc(kind=_lib.KIND_SIREN, D=len(self.network), W=self.hidden_dim, n_freq_xyz=0, n_freq_dir=0,
                       skip_mask=0, n_cls=0, logscale=1, box=2.0 / self.gridwarper.scale_factor, in_xyz=3, in_dir=3)

    def forward(self, input, z, ray_directions, **kwargs):
        # the reference's mapping network is commented out (models/nerf.py:185,198): same failure mode
        frequencies, phase_shifts = self.mapping_network(z)
        return self.forward_with_frequencies_phase_shifts(input, frequencies, phase_shifts, ray_directions, **kwargs)

    def forward_with_frequencies_phase_shifts(self, input, frequencies, phase_shifts, ray_directions, **kwargs):
        """input, ray_directions: (Bb, P, 3); frequencies, phase_shifts: (Bb, 9*hidden) -> (Bb, P, 4)."""
        if input.dim() != 3 or input.shape[-1] != 3:
            raise ValueError("input must be (batch, points, 3)")
        frequencies = frequencies * 15 + 30                                   # models/nerf.py:202
        outs = []
        for b in range(input.shape[0]):
            film = torch.cat([frequencies[b].reshape(-1), phase_shifts[b].reshape(-1)]).contiguous().float()
            x = torch.cat([input[b], ray_directions[b]], dim=-1)
            outs.append(_field.run_embedded(self, self._desc(), x, False, film=film))
        return torch.stack(outs, 0)


class SirenField(SemanticNeRF):
    """SIREN variant usable by render_rays: fixed FiLM vectors (buffers), ``forward(x[B,6], sigma_only)`` with
    x = [xyz, dir] (pair it with Embedding(3, 0) identity encodings).  frequencies = 0 gives the plain SIREN
    first-layer scale w0 = 30 (models/nerf.py:202)."""

    def __init__(self, hidden_dim=256, frequencies=None, phase_shifts=None):
        super().__init__(hidden_dim=hidden_dim)
        n = (len(self.network) + 1) * hidden_dim
        self.register_buffer("frequencies", torch.zeros(n) if frequencies is None else frequencies.reshape(-1).float())
        self.register_buffer("phase_shifts", torch.zeros(n) if phase_shifts is None else phase_shifts.reshape(-1).float())
        self.in_channels_xyz, self.in_channels_dir = 3, 3

    def _film(self):
        return torch.cat([self.frequencies * 15 + 30, self.phase_shifts]).contiguous()

    def forward(self, x, sigma_only=False):
        want = 3 if sigma_only else 6
        if x.dim() != 2 or x.shape[1] != want:
            raise ValueError(f"expected input of shape (B, {want}), got {tuple(x.shape)}")
        return _field.run_embedded(self, self._desc(), x, bool(sigma_only), film=self._film())
