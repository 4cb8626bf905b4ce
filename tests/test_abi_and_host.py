"""CPU-side checks: the C-ABI library loads and exports every symbol include/nerf_b200.h declares, the host
layer mirrors the reference's API surface (names, signatures, state_dict keys) and refuses CPU tensors loudly,
and the plain-C oracle agrees with the reference-derived fixtures.  No kernel is launched here."""
import ctypes as C
import inspect
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import nerf_siren_b200 as nb
    header = open(os.path.join(ROOT, "include", "nerf_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(nerfb_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 15
    raw = C.CDLL(nb._lib.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in the header but not exported"
    assert declared == set(nb._lib.SIGNATURES), declared ^ set(nb._lib.SIGNATURES)
    assert nb._lib.lib.nerfb_arch() == 100 and nb._lib.lib.nerfb_version() >= 100
    assert b"workspace" in nb._lib.lib.nerfb_error_string(-3)


def test_param_layout_matches_reference_state_dict():
    import nerf_siren_b200 as nb
    from oracle import render_oracle as O
    m = nb.NeRF()
    assert [(k, tuple(v.shape)) for k, v in m.state_dict().items()] == O.nerf_param_shapes()
    m3 = nb.NeRF_3D(CLS=6)
    assert [(k, tuple(v.shape)) for k, v in m3.state_dict().items()] == O.nerf_param_shapes(n_cls=6)
    for mod in (m, m3, nb.SirenField()):
        d = mod._desc(False)
        assert nb._lib.lib.nerfb_param_count(C.byref(d)) == sum(p.numel() for p in mod.parameters())
    s = nb.SemanticNeRF()
    assert set(s.state_dict()) == set(O.siren_init())
    assert nb._lib.lib.nerfb_out_channels(C.byref(m3._desc(False))) == 10
    bad = m._desc(False); bad.W = 7
    assert nb._lib.lib.nerfb_param_count(C.byref(bad)) == -1


def test_same_seed_gives_reference_default_init():
    """Registration order and default init consume the RNG exactly as nn.Linear stacks do in the reference."""
    import nerf_siren_b200 as nb
    torch.manual_seed(0)
    a = nb.NeRF()
    torch.manual_seed(0)
    lin = torch.nn.Linear(63, 256)
    assert torch.equal(a.state_dict()["xyz_encoding_1.0.weight"], lin.weight)


def test_signatures_match_reference_call_sites():
    import nerf_siren_b200 as nb
    sig = inspect.signature(nb.render_rays)
    assert list(sig.parameters) == ["models", "embeddings", "rays", "N_samples", "use_disp", "perturb", "noise_std",
                                    "N_importance", "chunk", "white_back", "test_time", "_cls_num", "network"]
    d = {k: v.default for k, v in sig.parameters.items() if v.default is not inspect._empty}
    assert d == dict(N_samples=64, use_disp=False, perturb=0, noise_std=1, N_importance=0, chunk=1024 * 32,
                     white_back=False, test_time=False, _cls_num=6, network=None)
    assert list(inspect.signature(nb.sample_pdf).parameters) == ["bins", "weights", "N_importance", "det", "eps"]
    e = nb.Embedding(3, 10)
    assert (e.N_freqs, e.in_channels, e.out_channels) == (10, 3, 63)
    assert torch.equal(e.freq_bands, 2.0 ** torch.arange(10))
    nb.install_as_models()
    from models.rendering import render_rays, render_rays_3d, render_rays_3d_conv  # noqa: F401  system.py:5
    from models.nerf import Embedding, NeRF  # noqa: F401                                       eval.py:10
    from models.nerf_cls import NeRF_3D  # noqa: F401


def test_cpu_tensors_are_refused_not_silently_computed():
    import nerf_siren_b200 as nb
    rays = torch.zeros(4, 8)
    with pytest.raises(RuntimeError, match="CUDA"):
        nb.render_rays([nb.NeRF()], [nb.Embedding(3, 10), nb.Embedding(3, 4)], rays)
    with pytest.raises(RuntimeError, match="CUDA"):
        nb.Embedding(3, 4)(torch.zeros(2, 3))
    with pytest.raises(RuntimeError, match="CUDA"):
        nb.NeRF()(torch.zeros(2, 90))
    with pytest.raises(ValueError):
        nb.render_rays([nb.NeRF()], [nb.Embedding(3, 10), nb.Embedding(3, 4)], torch.zeros(4, 7))


def test_c_oracle_matches_reference_fixtures():
    from tests._util import gold
    so = os.path.join(ROOT, "oracle", "liboracle_resample.so")
    if not os.path.exists(so):
        pytest.skip("oracle/liboracle_resample.so not built (run __graft_entry__.build())")
    lib = C.CDLL(so)
    g = gold("g8_searchsorted")
    a, v = np.ascontiguousarray(g["a"].numpy()), np.ascontiguousarray(g["v"].numpy())
    for side, left in (("left", 1), ("right", 0)):
        out = np.empty(v.shape, np.int64)
        lib.oracle_searchsorted(a.ctypes.data_as(C.c_void_p), C.c_int64(a.shape[0]), C.c_int64(a.shape[1]),
                                v.ctypes.data_as(C.c_void_p), C.c_int64(v.shape[0]), C.c_int64(v.shape[1]), left,
                                out.ctypes.data_as(C.c_void_p))
        assert np.array_equal(out, g[side].numpy())
    p = gold("g7_sample_pdf")
    for name in ("peaky", "flat", "zero", "rand"):
        cdf = np.ascontiguousarray(p[f"{name}_64_cdf"].numpy())
        u = np.ascontiguousarray(p[f"{name}_64_u"].numpy())
        inds, below, above = (np.empty(u.shape, np.int64) for _ in range(3))
        lib.oracle_pdf_indices(cdf.ctypes.data_as(C.c_void_p), u.ctypes.data_as(C.c_void_p), C.c_int64(u.shape[0]),
                               C.c_int64(cdf.shape[1]), C.c_int64(u.shape[1]), inds.ctypes.data_as(C.c_void_p),
                               below.ctypes.data_as(C.c_void_p), above.ctypes.data_as(C.c_void_p))
        assert np.array_equal(inds, p[f"{name}_64_inds_rnd"].numpy())
        assert below.min() >= 0 and above.max() <= 62
