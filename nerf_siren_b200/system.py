"""Lightning-free counterpart of the reference's ``NeRFSystem`` (system.py:172-306) and ``train.py:13-68`` (SURVEY 8f
row f1): the immediate caller of ``render_rays`` on the training side.  ``pytorch_lightning==1.2.0`` is absent on the
B200 box, so this module keeps the reference's method names and semantics (``forward`` chunk loop, ``training_step``,
``validation_step``, ``configure_optimizers``) and adds the small loop Lightning provided (``fit``), with checkpoints in
Lightning's key layout (``{'state_dict': {'nerf_coarse.<name>': ..., 'nerf_fine.<name>': ...}}``) so the reference's
``utils.load_ckpt(model, path, model_name='nerf_fine')`` (utils/__init__.py:56-86) reads them and vice versa.

hparams: any object with the fields of ``opt.py:3-87`` that this path uses: N_samples, N_importance, use_disp, perturb,
noise_std, chunk, batch_size, lr, optimizer ('adam' | 'sgd'), momentum, weight_decay, lr_scheduler ('steplr' | 'cosine' |
'poly'), decay_step, decay_gamma, poly_exp, num_epochs, loss_type ('mse').
"""
from __future__ import annotations

import os
from collections import defaultdict
from typing import Dict, Iterable, List, Optional

import torch
from torch import nn

from . import dp
from .losses import loss_dict
from .models.nerf import Embedding, NeRF
from .models.rendering import render_rays


# ---- utils/__init__.py:56-86 --------------------------------------------------------------------------------------
def extract_model_state_dict(ckpt_path, model_name="model", prefixes_to_ignore=()):
    checkpoint = torch.load(ckpt_path, map_location=torch.device("cpu"))
    if "state_dict" in checkpoint:                      # a pytorch-lightning checkpoint
        checkpoint = checkpoint["state_dict"]
    out = {}
    for k, v in checkpoint.items():
        if not k.startswith(model_name):
            continue
        k = k[len(model_name) + 1:]
        if any(k.startswith(p) for p in prefixes_to_ignore):
            continue
        out[k] = v
    return out


def load_ckpt(model: nn.Module, ckpt_path, model_name="model", prefixes_to_ignore=()):
    model_dict = model.state_dict()
    model_dict.update(extract_model_state_dict(ckpt_path, model_name, prefixes_to_ignore))
    model.load_state_dict(model_dict)


# ---- losses.py:10-20, metrics.py:4-14 (tiny tensor reductions on (B,3) outputs: host-side glue) -----------------------
def mse(image_pred, image_gt):
    return torch.mean((image_pred - image_gt) ** 2)


def psnr(image_pred, image_gt):
    return -10 * torch.log10(mse(image_pred, image_gt))


# ---- utils/__init__.py:11-52 --------------------------------------------------------------------------------------
def get_optimizer(hparams, models: Iterable[nn.Module]):
    params = [p for m in models for p in m.parameters()]
    name = getattr(hparams, "optimizer", "adam")
    if name == "sgd":
        return torch.optim.SGD(params, lr=hparams.lr, momentum=getattr(hparams, "momentum", 0.9),
                               weight_decay=getattr(hparams, "weight_decay", 0))
    if name == "adam":
        return torch.optim.Adam(params, lr=hparams.lr, eps=1e-8, weight_decay=getattr(hparams, "weight_decay", 0))
    raise ValueError("optimizer not recognized!")


def get_scheduler(hparams, optimizer):
    name = getattr(hparams, "lr_scheduler", "steplr")
    if name == "steplr":
        return torch.optim.lr_scheduler.MultiStepLR(optimizer, milestones=list(getattr(hparams, "decay_step", [20])),
                                                    gamma=getattr(hparams, "decay_gamma", 0.1))
    if name == "cosine":
        return torch.optim.lr_scheduler.CosineAnnealingLR(optimizer, T_max=hparams.num_epochs, eta_min=1e-8)
    if name == "poly":
        return torch.optim.lr_scheduler.LambdaLR(
            optimizer, lambda epoch: (1 - epoch / hparams.num_epochs) ** getattr(hparams, "poly_exp", 0.9))
    raise ValueError("scheduler not recognized!")


def get_learning_rate(optimizer):
    for g in optimizer.param_groups:
        return g["lr"]


class NeRFSystem(nn.Module):
    """system.py:172-306 without Lightning.  ``white_back`` replaces ``self.train_dataset.white_back`` (system.py:216)."""

    def __init__(self, hparams, white_back: bool = True, model_cls=NeRF, model_kwargs: Optional[dict] = None):
        super().__init__()
        self.hparams = hparams
        self.white_back = white_back
        self.embedding_xyz = Embedding(3, 10)
        self.embedding_dir = Embedding(3, 4)
        self.embeddings = [self.embedding_xyz, self.embedding_dir]
        kw = model_kwargs or {}
        self.nerf_coarse = model_cls(**kw)
        self.models = [self.nerf_coarse]
        if hparams.N_importance > 0:
            self.nerf_fine = model_cls(**kw)
            self.models += [self.nerf_fine]
        if getattr(hparams, "pretrained", None):
            load_ckpt(self.nerf_coarse, hparams.pretrained, model_name="nerf_coarse")
            if hparams.N_importance > 0:
                load_ckpt(self.nerf_fine, hparams.pretrained, model_name="nerf_fine")
        self.global_step, self.current_epoch = 0, 0
        self.optimizer = self.scheduler = None

    def decode_batch(self, batch):
        return batch["rays"], batch["rgbs"]

    def forward(self, rays) -> Dict[str, torch.Tensor]:
        """system.py:199-223: batched inference on rays using chunk (10 positional args, as the reference calls it)."""
        rays = rays.reshape(-1, 8)
        results = defaultdict(list)
        hp = self.hparams
        for i in range(0, rays.shape[0], hp.chunk):
            out = render_rays(self.models, self.embeddings, rays[i:i + hp.chunk], hp.N_samples, hp.use_disp, hp.perturb,
                              hp.noise_std, hp.N_importance, hp.chunk, self.white_back)
            for k, v in out.items():
                results[k] += [v]
        return {k: torch.cat(v, 0) for k, v in results.items()}

    def configure_optimizers(self):
        self.optimizer = get_optimizer(self.hparams, self.models)
        self.scheduler = get_scheduler(self.hparams, self.optimizer)
        return [self.optimizer], [self.scheduler]

    def training_step(self, batch, batch_nb=0):
        log = {"lr": get_learning_rate(self.optimizer)}
        rays, rgbs = self.decode_batch(batch)
        results = self(rays)
        log["train/loss"] = loss = dp.mse_loss(results, rgbs)
        typ = "fine" if "rgb_fine" in results else "coarse"
        with torch.no_grad():
            log["train/psnr"] = psnr_ = psnr(results[f"rgb_{typ}"], rgbs)
        return {"loss": loss, "progress_bar": {"train_psnr": psnr_}, "log": log}

    @torch.no_grad()
    def validation_step(self, batch, batch_nb=0):
        rays, rgbs = self.decode_batch(batch)
        rays, rgbs = rays.squeeze(), rgbs.squeeze()
        results = self(rays)
        typ = "fine" if "rgb_fine" in results else "coarse"
        return {"val_loss": dp.mse_loss(results, rgbs), "val_psnr": psnr(results[f"rgb_{typ}"], rgbs)}

    def validation_epoch_end(self, outputs):
        mean_loss = torch.stack([x["val_loss"] for x in outputs]).mean()
        mean_psnr = torch.stack([x["val_psnr"] for x in outputs]).mean()
        return {"progress_bar": {"val_loss": mean_loss, "val_psnr": mean_psnr},
                "log": {"val/loss": mean_loss, "val/psnr": mean_psnr}}

    # ---- what Lightning's Trainer did around these hooks (train.py:41-66) ----
    def state_dict_lightning(self):
        sd = {f"nerf_coarse.{k}": v.detach().cpu() for k, v in self.nerf_coarse.state_dict().items()}
        if len(self.models) > 1:
            sd.update({f"nerf_fine.{k}": v.detach().cpu() for k, v in self.nerf_fine.state_dict().items()})
        return sd

    def save_checkpoint(self, path):
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        torch.save({"epoch": self.current_epoch, "global_step": self.global_step, "state_dict": self.state_dict_lightning(),
                    "optimizer_states": [self.optimizer.state_dict()] if self.optimizer is not None else [],
                    "lr_schedulers": [self.scheduler.state_dict()] if self.scheduler is not None else []}, path)

    def fit(self, train_batches: Iterable[dict], val_batches: Optional[Iterable[dict]] = None, ckpt_dir: Optional[str] = None,
            log_every: int = 0) -> List[dict]:
        """One optimizer / scheduler pair, ``num_epochs`` passes over ``train_batches`` (an iterable of
        {'rays': (B,8), 'rgbs': (B,3)} CUDA batches, re-iterable), scheduler.step() once per epoch as Lightning does,
        validation + checkpoint at each epoch end.  Returns the per-epoch logs."""
        if self.optimizer is None:
            self.configure_optimizers()
        history = []
        for epoch in range(self.hparams.num_epochs):
            self.current_epoch = epoch
            last = None
            for nb_, batch in enumerate(train_batches):
                self.optimizer.zero_grad(set_to_none=True)
                out = self.training_step(batch, nb_)
                out["loss"].backward()
                self.optimizer.step()
                self.global_step += 1
                last = out["log"]
                if log_every and self.global_step % log_every == 0:
                    print(f"epoch {epoch} step {self.global_step} loss {float(last['train/loss']):.5f} "
                          f"psnr {float(last['train/psnr']):.2f} lr {last['lr']:.2e}", flush=True)
            self.scheduler.step()
            log = {"epoch": epoch, **{k: float(v.detach() if torch.is_tensor(v) else v) for k, v in (last or {}).items()}}
            if val_batches is not None:
                outs = [self.validation_step(b, i) for i, b in enumerate(val_batches)]
                log.update({k: float(v) for k, v in self.validation_epoch_end(outs)["log"].items()})
            if ckpt_dir:
                self.save_checkpoint(os.path.join(ckpt_dir, f"epoch={epoch}.ckpt"))
            history.append(log)
        return history


class EG3DSystem(nn.Module):
    """system.py:17-170 without Lightning: the caller of the EG3D path on the training side (``train.py`` builds it for the
    tri-plane runs).  ``forward`` renders a batch of (B, 8) rays through ``EG3D_Renderer.render`` with ``conditioning_params =
    -1`` (system.py:72-75); ``training_step`` applies ``loss_dict[hparams.loss_type]`` to the six-key result and back-propagates
    through the hand-written kernels (fused render keeping its samples -> samples backward -> field backward);
    ``validation_step`` renders in chunks of 4096 rays (system.py:139-146).  ``plane_source`` / ``rendering_kwargs`` are passed
    on to ``EG3D_Renderer`` (the reference's StyleGAN2 backbone is outside this path; default: directly optimised planes)."""

    def __init__(self, hparams, plane_source=None, rendering_kwargs=None):
        super().__init__()
        from .eg3d import EG3D_Renderer
        self.hparams = hparams
        self.loss = loss_dict[getattr(hparams, "loss_type", "mse")]()
        self.eg3d_renderer = EG3D_Renderer(plane_source, rendering_kwargs)
        self.models = [self.eg3d_renderer]
        if getattr(hparams, "pretrained", None):
            load_ckpt(self.eg3d_renderer, hparams.pretrained, model_name="eg3d_renderer")
        self.global_step, self.current_epoch = 0, 0
        self.optimizer = self.scheduler = None

    def decode_batch(self, batch):
        return batch["rays"], batch["rgbs"]

    def forward(self, rays) -> Dict[str, torch.Tensor]:
        rays = rays.reshape(-1, 8)
        return self.eg3d_renderer.render(-1, rays[:, :3].contiguous(), rays[:, 3:6].contiguous())

    def configure_optimizers(self):
        self.optimizer = get_optimizer(self.hparams, self.models)
        self.scheduler = get_scheduler(self.hparams, self.optimizer)
        return [self.optimizer], [self.scheduler]

    def training_step(self, batch, batch_nb=0):
        log = {"lr": get_learning_rate(self.optimizer)}
        rays, rgbs = self.decode_batch(batch)
        results = self(rays)
        log["train/loss"] = loss = self.loss(results, rgbs)
        typ = "fine" if "rgb_fine" in results else "coarse"
        with torch.no_grad():
            log["train/psnr"] = psnr_ = psnr(results[f"rgb_{typ}"], rgbs)
        return {"loss": loss, "progress_bar": {"train_psnr": psnr_}, "log": log}

    @torch.no_grad()
    def validation_step(self, batch, batch_nb=0):
        rays, rgbs = self.decode_batch(batch)
        rays, rgbs = rays.squeeze(), rgbs.squeeze()
        chunk = 1024 * 4
        results = defaultdict(list)
        for i in range(0, rays.shape[0], chunk):
            for k, v in self(rays[i:i + chunk]).items():
                results[k] += [v]
        results = {k: torch.cat(v, 0) for k, v in results.items()}
        typ = "fine" if "rgb_fine" in results else "coarse"
        return {"val_loss": self.loss(results, rgbs), "val_psnr": psnr(results[f"rgb_{typ}"], rgbs)}

    validation_epoch_end = NeRFSystem.validation_epoch_end

    def state_dict_lightning(self):
        return {f"eg3d_renderer.{k}": v.detach().cpu() for k, v in self.eg3d_renderer.state_dict().items()}

    save_checkpoint = NeRFSystem.save_checkpoint
    fit = NeRFSystem.fit
