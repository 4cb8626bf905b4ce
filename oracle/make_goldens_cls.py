"""G11: the semantic-head training boundary of BASELINE configs[3] -- class compositing
(models/rendering.py:379) + the reference's own ``MSECELoss`` / ``MSENLLLoss`` (losses.py:23-100) + autograd through
the reference's own ``NeRF_3D`` modules (models/nerf_cls.py:12-112).

Runs only in the build container (imports /root/reference read-only; never copies it).  ``render_rays_3d``
itself hard-codes ``.cuda()`` and the out-of-scope PointNet branch (rendering.py:352), so the render around the
reference modules is the oracle's ``render_rays`` -- which oracle/make_goldens.py pins bit-exact against the
reference's ``render_rays`` -- with the class channels composited per rendering.py:379.  Everything else on the path
of the stored numbers is reference code: the two NeRF_3D networks, the loss modules, torch autograd.

    python oracle/make_goldens_cls.py        # writes tests/golden/g11_cls_loss.npz

TEST INFRASTRUCTURE.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import render_oracle as O                    # noqa: E402
from oracle.make_goldens import GOLD, close, grad_slices, import_reference, np_   # noqa: E402


@contextlib.contextmanager
def _cuda_is_identity():
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        yield
    finally:
        torch.Tensor.cuda = orig


def main():
    ref_nerf, ref_rend, ref_cls = import_reference()
    import losses as ref_losses                            # /root/reference/losses.py
    assert ref_cls is not None, "models.nerf_cls must import (MinkowskiEngine / open3d stubs)"
    torch.set_num_threads(1)
    g = torch.Generator().manual_seed(4321)
    n, sc, si, n_cls = 48, 64, 64, 6
    rays = O.llff_rays(n, seed=2)
    pc, pf = O.nerf_field(seed=31, n_cls=n_cls), O.nerf_field(seed=32, n_cls=n_cls)
    rng = dict(perturb_rand=torch.rand(n, sc, generator=g), noise_coarse=torch.randn(n, sc, generator=g),
               u=torch.rand(n, si, generator=g), noise_fine=torch.randn(n, sc + si, generator=g))
    target = torch.rand(n, 3, generator=g)
    labels = torch.randint(0, n_cls, (n, 1), generator=g)
    labels[::7] = -1                                       # ignore_index rows (losses.py:27)
    save = {"rays": rays, "target": target, "labels": labels}
    save.update({"rng:" + k: v for k, v in rng.items()})

    def run(loss_mod, weight, log_softmax, tag, labels_used):
        mc, mf = ref_cls.NeRF_3D(CLS=n_cls), ref_cls.NeRF_3D(CLS=n_cls)
        mc.load_state_dict(pc); mf.load_state_dict(pf)
        fields = [lambda x, so: mc(x, sigma_only=so), lambda x, so: mf(x, sigma_only=so)]
        out = O.render_rays(fields, rays, sc, False, 1.0, 1.0, si, 32768, False, False, rng=rng)
        res = {k: v for k, v in out.items() if not k.startswith("_")}
        if log_softmax:                                    # render_rays_3d_conv, rendering.py:651-652
            for k in ("cls_coarse", "cls_fine"):
                res[k] = torch.log_softmax(res[k], dim=-1)
        # losses.py:58 prints; losses.py:75,90 call .cuda() on the class maps (no GPU here: identity for the call)
        with contextlib.redirect_stdout(io.StringIO()), _cuda_is_identity():
            if weight is None:
                loss = loss_mod(res, target, labels_used)  # exactly as system.py:351 calls it
            else:
                loss = loss_mod(res, target, labels_used, weight)
        loss["sum"].backward()
        # the oracle's own parameter-dict field must agree with the reference modules it stands in for
        pco = {k: v.clone().requires_grad_(True) for k, v in pc.items()}
        pfo = {k: v.clone().requires_grad_(True) for k, v in pf.items()}
        fo = [lambda x, so: O.nerf_mlp(pco, x, so), lambda x, so: O.nerf_mlp(pfo, x, so)]
        oo = O.render_rays(fo, rays, sc, False, 1.0, 1.0, si, 32768, False, False, rng=rng)
        for k in ("cls_coarse", "cls_fine", "rgb_coarse", "rgb_fine"):
            assert torch.equal(oo[k], out[k]), k
        save.update({f"{tag}:loss_{k}": v.detach() for k, v in loss.items()})
        save.update({f"{tag}:gradc:" + k: v for k, v in grad_slices({k: p.grad.clone() for k, p in mc.named_parameters()}).items()})
        save.update({f"{tag}:gradf:" + k: v for k, v in grad_slices({k: p.grad.clone() for k, p in mf.named_parameters()}).items()})
        if tag == "ce_default":
            save.update({"out:" + k: v.detach() for k, v in res.items()})
        return loss

    run(ref_losses.MSECELoss(), None, False, "ce_default", labels)       # weight = 0. (the reference's default)
    run(ref_losses.MSECELoss(), 0.5, False, "ce_half", labels)
    nll_labels = labels.clamp_min(0)                                      # nll_loss has no ignore_index here (:84)
    save["labels_nll"] = nll_labels
    run(ref_losses.MSENLLLoss(), None, True, "nll", nll_labels)
    np.savez_compressed(os.path.join(GOLD, "g11_cls_loss.npz"), **np_(save))
    print("wrote g11_cls_loss.npz:", {k: float(v) for k, v in save.items() if ":loss_" in k})


if __name__ == "__main__":
    main()
