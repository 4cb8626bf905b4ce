"""SURVEY 8e on CPU: the N > 1 host logic of the ray-sharded data-parallel step with world_size = 2 over gloo.
What runs here is the product's own code up to the collective -- `shard_bounds`, `FlatParameters` (one flat parameter /
gradient buffer whose views the modules keep), `DataParallelStep.forward_backward` and `allreduce_sum_` -- around a small
differentiable CPU stand-in for `render_rays` (the kernels and the Adam launch need a GPU and are covered by `-m gpu`)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _Field(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(8, 16)
        self.b = torch.nn.Linear(16, 3)

    def forward(self, rays):
        return torch.sigmoid(self.b(torch.relu(self.a(rays))))


def _render_stub(models, embeddings, rays, *args):
    out = {"rgb_coarse": models[0](rays)}
    if len(models) > 1:
        out["rgb_fine"] = models[1](rays)
    return out


def _make_models():
    torch.manual_seed(7)                       # identical init on every rank, as DDP's broadcast would give
    return [_Field(), _Field()]


def _batch(n):
    g = torch.Generator().manual_seed(11)
    return torch.randn(n, 8, generator=g), torch.rand(n, 3, generator=g)


def _worker(rank, world, port, n_rays, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nerf_siren_b200 import dp
        rays, rgbs = _batch(n_rays)
        lo, hi = dp.shard_bounds(n_rays, rank, world)
        step = dp.DataParallelStep(_make_models(), [None, None], render_fn=_render_stub)
        loss = step.forward_backward(rays[lo:hi], rgbs[lo:hi])
        local = step.params.flat.grad.clone()
        dp.allreduce_sum_(step.params.flat.grad, step.group)
        # the modules' .grad are views of the flat buffer: they see the reduced values without a copy
        assert step.models[0].a.weight.grad.data_ptr() == step.params.flat.grad.data_ptr()
        torch.save({"lo": lo, "hi": hi, "loss": loss, "local": local, "reduced": step.params.flat.grad.clone(),
                    "world": dp.world_size(step.group)}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_rays", [64, 37])
def test_two_rank_gradient_allreduce_matches_single_process(tmp_path, n_rays):
    pytest.importorskip("nerf_siren_b200")
    from nerf_siren_b200 import dp
    world = 2
    mp.get_context("spawn")
    mp.spawn(_worker, args=(world, _free_port(), n_rays, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(os.path.join(tmp_path, f"rank{k}.pt")) for k in range(world)]
    # shards: contiguous, disjoint, covering (ragged when n_rays is odd)
    assert r[0]["lo"] == 0 and r[0]["hi"] == r[1]["lo"] and r[1]["hi"] == n_rays and all(x["world"] == 2 for x in r)
    # the collective: both ranks end with the bit-identical SUM of the two local gradients
    assert torch.equal(r[0]["reduced"], r[1]["reduced"])
    assert torch.allclose(r[0]["reduced"], r[0]["local"] + r[1]["local"], rtol=0, atol=1e-7)
    # DDP semantics (mean over ranks of each rank's mean loss): the 1/world scale the Adam kernel applies to the SUM
    rays, rgbs = _batch(n_rays)
    want = torch.zeros_like(r[0]["reduced"])
    for k in range(world):
        step = dp.DataParallelStep(_make_models(), [None, None], render_fn=_render_stub)
        step.forward_backward(rays[r[k]["lo"]:r[k]["hi"]], rgbs[r[k]["lo"]:r[k]["hi"]])
        want += step.params.flat.grad / world
        assert torch.allclose(step.params.flat.grad, r[k]["local"], rtol=0, atol=1e-7)
    assert torch.allclose(r[0]["reduced"] / world, want, rtol=0, atol=1e-7)


def test_shard_bounds_partition():
    pytest.importorskip("nerf_siren_b200")
    from nerf_siren_b200 import dp
    for n in (0, 1, 7, 640000):
        for world in (1, 2, 3, 8):
            b = [dp.shard_bounds(n, k, world) for k in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[k][1] == b[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_loss_hook_receives_labels_and_dict_losses_backprop_their_sum():
    """configs[3] host logic: DataParallelStep(loss_fn=MSECELoss()) calls loss_fn(results, rgbs, labels) as system.py:351
    does and back-propagates the dict's "sum"; render_kwargs reach the render function."""
    nb = pytest.importorskip("nerf_siren_b200")
    from nerf_siren_b200 import dp, losses
    seen = {}

    class _ClsField(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.rgb, self.cls = torch.nn.Linear(8, 3), torch.nn.Linear(8, 6)

    def render(models, embeddings, rays, *args, **kw):
        seen.update(kw)
        out = {}
        for m, t in zip(models, ("coarse", "fine")):
            out["rgb_" + t], out["cls_" + t] = torch.sigmoid(m.rgb(rays)), m.cls(rays)
        return out
    torch.manual_seed(3)
    step = dp.DataParallelStep([_ClsField(), _ClsField()], [None, None], render_fn=render, loss_fn=losses.MSECELoss(),
                               render_kwargs={"no_grad_on_nerf": False}, loss_kwargs={"weight": 0.25})
    rays, rgbs = _batch(40)
    labels = torch.randint(0, 6, (40, 1), generator=torch.Generator().manual_seed(5))
    loss = step.forward_backward(rays, rgbs, labels)
    assert seen == {"no_grad_on_nerf": False}
    want = losses.MSECELoss()(render(step.models, None, rays), rgbs, labels, weight=0.25)["sum"]
    assert float(loss) == pytest.approx(float(want), rel=1e-6)
    assert all(float(m.cls.weight.grad.abs().sum()) > 0 and float(m.rgb.weight.grad.abs().sum()) > 0 for m in step.models)
