"""Host-side loss glue (losses.py:10-100 of the reference) against golden G11, which holds the values the REFERENCE's
own MSECELoss / MSENLLLoss returned on reference-composited class maps (oracle/make_goldens_cls.py).  CPU only."""
import pytest
import torch

from tests._util import gold


def _outs(g):
    return {k[4:]: v.clone() for k, v in g.items() if k.startswith("out:")}


def test_msece_and_msenll_match_reference_values():
    nb = pytest.importorskip("nerf_siren_b200")
    g = gold("g11_cls_loss")
    res = _outs(g)
    assert {"rgb_coarse", "rgb_fine", "cls_coarse", "cls_fine"} <= set(res)
    for tag, args in (("ce_default", ()), ("ce_half", (0.5,))):
        loss = nb.losses.MSECELoss()(res, g["target"], g["labels"], *args)
        for k in ("sum", "rgb", "cls"):
            assert float(loss[k]) == pytest.approx(float(g[f"{tag}:loss_{k}"]), rel=1e-6, abs=1e-9), (tag, k)
    assert float(nb.losses.MSECELoss()(res, g["target"], g["labels"])["rgb"]) == 0.0      # weight = 0. default (losses.py:29)
    res_log = dict(res)
    for k in ("cls_coarse", "cls_fine"):
        res_log[k] = torch.log_softmax(res[k], dim=-1)                                     # rendering.py:651-652
    loss = nb.losses.MSENLLLoss()(res_log, g["target"], g["labels_nll"], 0.3)               # the weight argument is ignored (:95)
    for k in ("sum", "rgb", "cls"):
        assert float(loss[k]) == pytest.approx(float(g[f"nll:loss_{k}"]), rel=1e-6, abs=1e-9), k


def test_loss_dict_and_mse():
    nb = pytest.importorskip("nerf_siren_b200")
    assert set(nb.losses.loss_dict) == {"mse", "msece", "msenll"}
    g = gold("g9_blender_train")
    res = {k[4:]: v for k, v in g.items() if k.startswith("out:")}
    assert float(nb.losses.loss_dict["mse"]()(res, g["target"])) == pytest.approx(float(g["loss"]), rel=1e-6)
    nb.install_as_losses()
    from losses import loss_dict                      # system.py:10
    assert loss_dict is nb.losses.loss_dict


def test_ignore_index_rows_carry_no_class_gradient():
    nb = pytest.importorskip("nerf_siren_b200")
    g = gold("g11_cls_loss")
    res = _outs(g)
    res["cls_fine"].requires_grad_(True)
    nb.losses.MSECELoss()(res, g["target"], g["labels"])["sum"].backward()
    ignored = (g["labels"].reshape(-1) == -1)
    assert ignored.any() and float(res["cls_fine"].grad[ignored].abs().sum()) == 0.0
    assert float(res["cls_fine"].grad[~ignored].abs().sum()) > 0.0
