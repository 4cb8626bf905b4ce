"""Step (5): the ray-sharded data-parallel training step.

Reference: one Lightning process per GPU under DDP (train.py:41-66), ``batch_size`` rays PER RANK
(system.py:243-248), ``MSELoss`` on rgb_coarse + rgb_fine (losses.py:10-20, system.py:257-275), DDP's bucketed
gradient mean, then ``torch.optim.Adam(lr=5e-4, eps=1e-8)`` (utils/__init__.py:11-31).

Here: one process per GPU over ``torch.distributed`` (NCCL on NVLink/NVSwitch; gloo in the CPU tests).  All
parameters of [coarse, fine] are re-homed into ONE contiguous fp32 buffer (the modules keep their reference
parameter names -- each ``nn.Parameter`` becomes a view), gradients land in one flat buffer, and the only
collective of the step is ONE all-reduce (SUM) of that buffer at the batch boundary; the 1/world_size of the DDP
mean is folded into the single-launch Adam kernel (``nerfb_adam_step``).  No activation, ray or sample is ever
exchanged: rays are independent, every rank renders its own shard.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.distributed as dist
from torch import nn

from . import ops


def shard_bounds(n: int, rank: int, world: int):
    """Contiguous block partition of n rays over `world` ranks (full-image rendering): rank r owns
    [lo, hi); the first n % world ranks get one extra ray."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_sum_(flat: torch.Tensor, group=None) -> torch.Tensor:
    """The one collective of the training step: in-place SUM of the flat gradient buffer across ranks.  A no-op
    without an initialised process group (single GPU)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def world_size(group=None) -> int:
    return dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1


class FlatParameters:
    """Moves the parameters of several modules into one contiguous fp32 buffer.

    ``flat`` is the autograd leaf; every module parameter keeps its name and shape but its storage becomes a view
    of ``flat`` (so ``state_dict`` / ``load_state_dict`` / ``utils.load_ckpt`` prefixes keep working), and its
    ``.grad`` a view of ``flat.grad``.  The fused render path picks the flat slice up through
    ``module._nerfb_flat_slice`` instead of concatenating parameters every call.
    """

    def __init__(self, modules: Sequence[nn.Module]):
        params = [p for m in modules for p in m.parameters()]
        if not params:
            raise ValueError("no parameters to flatten")
        dev = params[0].device
        if any(p.device != dev or p.dtype != torch.float32 for p in params):
            raise ValueError("all parameters must be float32 on one device")
        total = sum(p.numel() for p in params)
        # every model's slice starts 16-byte aligned (float4 Adam kernel); NeRF sizes are multiples of 4 already
        self.flat = torch.empty(total, device=dev, dtype=torch.float32)
        self.flat.requires_grad_(True)
        self.flat.grad = torch.zeros(total, device=dev, dtype=torch.float32)
        self.slices = []
        off = 0
        for m in modules:
            start = off
            for p in m.parameters():
                n = p.numel()
                with torch.no_grad():
                    self.flat.data[off:off + n].copy_(p.data.reshape(-1))
                p.data = self.flat.data[off:off + n].view_as(p)
                p.grad = self.flat.grad[off:off + n].view_as(p)
                off += n
            self.slices.append((start, off))
            m._nerfb_flat_slice = _FlatSlice(self, start, off, m)

    def zero_grad(self):
        self.flat.grad.zero_()


class _FlatSlice:
    def __init__(self, owner: FlatParameters, lo: int, hi: int, module: nn.Module):
        self.owner, self.lo, self.hi = owner, lo, hi
        self._ptrs = [p.data_ptr() for p in module.parameters()]

    def get(self, module: nn.Module) -> Optional[torch.Tensor]:
        """The differentiable flat view, or None when the module's storage was re-homed since (``.to()``...)."""
        if [p.data_ptr() for p in module.parameters()] != self._ptrs:
            return None
        return self.owner.flat[self.lo:self.hi]


def mse_loss(results: dict, rgbs: torch.Tensor) -> torch.Tensor:
    """losses.py:15-20: mean((rgb_coarse - t)^2) (+ mean((rgb_fine - t)^2) when present)."""
    loss = ((results["rgb_coarse"] - rgbs) ** 2).mean()
    if "rgb_fine" in results:
        loss = loss + ((results["rgb_fine"] - rgbs) ** 2).mean()
    return loss


class DataParallelStep:
    """render_rays -> loss -> backward -> ONE flat all-reduce -> Adam, for this rank's shard of rays.

    ``loss_fn(results, rgbs)`` defaults to MSELoss (losses.py:10-20).  For the semantic head (BASELINE configs[3]) pass
    ``render_fn=render_rays_3d, render_kwargs={"no_grad_on_nerf": False}`` and ``loss_fn=losses.MSECELoss()`` /
    ``MSENLLLoss()``: ``step(rays, rgbs, labels)`` then calls ``loss_fn(results, rgbs, labels)`` the way
    system.py:351 does and back-propagates its ``"sum"`` entry."""

    def __init__(self, models: List[nn.Module], embeddings: List[nn.Module], N_samples=64, N_importance=64,
                 use_disp=False, perturb=1.0, noise_std=1.0, chunk=32 * 1024, white_back=True, lr=5e-4,
                 betas=(0.9, 0.999), eps=1e-8, group=None, render_fn=None, cuda_graph=False, loss_fn=None,
                 render_kwargs=None, loss_kwargs=None):
        from .models.rendering import render_rays
        self.models, self.embeddings = list(models), list(embeddings)
        self.render_fn = render_fn or render_rays
        self.render_args = (N_samples, use_disp, perturb, noise_std, N_importance, chunk, white_back)
        self.render_kwargs = dict(render_kwargs or {})
        self.loss_fn, self.loss_kwargs = loss_fn, dict(loss_kwargs or {})
        self.group = group
        self.lr, self.betas, self.eps = lr, betas, eps
        self.params = FlatParameters(self.models)
        self.exp_avg = torch.zeros_like(self.params.flat.data)
        self.exp_avg_sq = torch.zeros_like(self.params.flat.data)
        self.steps = 0
        # cuda_graph: forward + backward of one batch shape is captured once and replayed (the ~60 launches of a step
        # cost more host time than the GPU needs for the small ones); the all-reduce and the Adam kernel stay outside,
        # so the step counter and the collective are ordinary launches.
        self.cuda_graph = bool(cuda_graph)
        self._graphs = {}
        self.graph_launches = 0            # kernels of this library launched by graph replays (ops.launch_count() only sees host calls)

    def _captured(self, rays: torch.Tensor, rgbs: torch.Tensor, labels: Optional[torch.Tensor]):
        key = (tuple(rays.shape), tuple(rgbs.shape), None if labels is None else tuple(labels.shape))
        st = self._graphs.get(key)
        if st is None:
            import types
            dev = self.params.flat.device
            st = types.SimpleNamespace(rays=torch.empty(rays.shape, device=dev, dtype=torch.float32),
                                       rgbs=torch.empty(rgbs.shape, device=dev, dtype=torch.float32),
                                       labels=None if labels is None else torch.empty(labels.shape, device=dev, dtype=labels.dtype))
            st.rays.copy_(rays); st.rgbs.copy_(rgbs)
            if labels is not None:
                st.labels.copy_(labels)
            side = torch.cuda.Stream(device=dev)             # warm-up off the capture stream (allocator, autograd, lazy inits)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(2):
                    self.forward_backward(st.rays, st.rgbs, st.labels)
            torch.cuda.current_stream(dev).wait_stream(side)
            before = ops.launch_count()
            st.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(st.graph):
                st.loss = self.forward_backward(st.rays, st.rgbs, st.labels)
            st.launches = ops.launch_count() - before
            self._graphs[key] = st
        return st

    def forward_backward(self, rays: torch.Tensor, rgbs: torch.Tensor, labels: Optional[torch.Tensor] = None) -> torch.Tensor:
        self.params.zero_grad()
        results = self.render_fn(self.models, self.embeddings, rays, *self.render_args, **self.render_kwargs)
        if self.loss_fn is None:
            loss = mse_loss(results, rgbs)
        elif labels is None:
            loss = self.loss_fn(results, rgbs, **self.loss_kwargs)
        else:
            loss = self.loss_fn(results, rgbs, labels, **self.loss_kwargs)
        if isinstance(loss, dict):                            # MSECELoss / MSENLLLoss: {"sum", "rgb", "cls"}, system.py:352
            loss = loss["sum"]
        loss.backward()
        return loss.detach()

    def step(self, rays: torch.Tensor, rgbs: torch.Tensor, labels: Optional[torch.Tensor] = None) -> torch.Tensor:
        if self.cuda_graph and self.params.flat.is_cuda:
            st = self._captured(rays, rgbs, labels)
            st.rays.copy_(rays, non_blocking=True)          # host (pinned) or device batch -> the graph's input buffers
            st.rgbs.copy_(rgbs, non_blocking=True)
            if labels is not None:
                st.labels.copy_(labels, non_blocking=True)
            st.graph.replay()
            self.graph_launches += st.launches
            loss = st.loss.clone()
        else:
            loss = self.forward_backward(rays, rgbs, labels)
        g = self.params.flat.grad
        allreduce_sum_(g, self.group)                       # the batch-boundary collective (4.77 MB for NeRF x2)
        self.steps += 1
        ops.adam_step(self.params.flat.data, g, self.exp_avg, self.exp_avg_sq, self.steps, self.lr, self.betas,
                      self.eps, grad_scale=1.0 / world_size(self.group))
        return loss
