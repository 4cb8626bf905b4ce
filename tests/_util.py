"""Shared helpers for the test-suite: golden loading and oracle-side model closures."""
import os

import numpy as np
import torch

from oracle import render_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gold(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: torch.from_numpy(np.asarray(z[k])) for k in z.files}


def checksum(p):
    return float(sum(v.double().abs().sum() for v in p.values()))


def fields(pc, pf):
    return [lambda x, so, p=pc: O.nerf_mlp(p, x, so), lambda x, so, p=pf: O.nerf_mlp(p, x, so)]


def rng_of(g):
    return {k[4:]: v for k, v in g.items() if k.startswith("rng:")}


# the render configurations of oracle/make_goldens.py (G9), keyed by fixture name
BASE = dict(n_samples=64, use_disp=False, perturb=0.0, noise_std=0.0, n_importance=64, chunk=32768,
            white_back=True, test_time=False)
CASES = {
    "blender_det": dict(BASE),
    "blender_train": dict(BASE, perturb=1.0, noise_std=1.0),
    "blender_eval": dict(BASE, test_time=True),
    "blender_eval128": dict(BASE, test_time=True, n_importance=128),
    "blender_coarse_only": dict(BASE, n_importance=0),
    "blender_disp": dict(BASE, use_disp=True, perturb=1.0),
    "llff_train": dict(BASE, perturb=1.0, noise_std=1.0, white_back=False),
    "llff_cls": dict(BASE, perturb=1.0, noise_std=1.0, white_back=False),
}


def case_params(name):
    if name == "llff_cls":
        return O.nerf_field(seed=31, n_cls=6), O.nerf_field(seed=32, n_cls=6)
    return O.nerf_field(seed=21), O.nerf_field(seed=22)
