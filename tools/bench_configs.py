"""The other BASELINE.json configurations, one JSON line each (bench.py itself measures configs[1], the headline):

  render   NeRF coarse 64 + fine 64 full-image render (test_time=True), 800 x 800 Blender-shaped rays      [metric "render"]
  siren    configs[2]: SIREN variant render, 64 + 128 samples; one GPU's shard of an 800 x 800 image under 8-way ray sharding
           (80,000 rays) -- or the shard for --gpus N under torchrun
  cls      configs[3]: NeRF_3D (CLS = 6) semantic-head training step, LLFF-shaped rays, batch 8192 per GPU
  eg3d     configs[4]: see tools/bench_eg3d.py

Timing: CUDA events around whole public-API calls (render_rays / DataParallelStep.step) on resident inputs, after warm-up,
max over ranks when launched under torchrun.  MLP FLOPs per sample from SURVEY 8d.
    python tools/bench_configs.py [render|siren|cls|siren_train|render_e2e ...]
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_siren_b200 as nb  # noqa: E402
from nerf_siren_b200 import synth  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
PEAK = peaks.get("bf16_tflops_sustained", 1400.0)


def timed(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def emit(line):
    if rank == 0:
        print(json.dumps(line), flush=True)


def render(kind):
    torch.manual_seed(0)
    nb.set_precision("auto")
    if kind == "siren":
        models = [nb.SirenField().to(dev), nb.SirenField().to(dev)]
        emb = [nb.Embedding(3, 0), nb.Embedding(3, 0)]
        S, Si, flop_c, flop_f = 64, 128, 2 * 459_776, 2 * 526_848            # sigma-only coarse under test_time
    else:
        models = [nb.NeRF().to(dev), nb.NeRF().to(dev)]
        emb = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
        S, Si, flop_c, flop_f = 64, 64, 2 * 491_264, 2 * 593_408
    rays = synth.blender_image_rays(800, 800, seed=1)
    n_img = rays.shape[0]
    shard = 8 if (kind == "siren" and world == 1) else world               # config 3 is quoted on 8-way ray sharding
    lo, hi = nb.dp.shard_bounds(n_img, rank if world > 1 else 0, shard)
    rays = rays[lo:hi].contiguous().to(dev)
    n = rays.shape[0]

    def run():
        with torch.no_grad():
            return nb.render_rays(models, emb, rays, S, False, 0, 0, Si, 32768, True, test_time=True)
    ms = timed(run, 3)
    flops = n * (S * flop_c + (S + Si) * flop_f)
    emit({"metric": f"rays/s, {'SIREN' if kind == 'siren' else 'NeRF'} full-image render ({S}+{Si} samples, test_time)",
          "value": n * world / ms * 1e3, "unit": "rays/s", "n_gpus": world, "rays_per_gpu": n, "ms": ms,
          "mlp_tflops": flops / ms / 1e9, "frac_of_sustained_bf16_peak": flops / ms / 1e9 / PEAK, "dtype": "bf16",
          "config": {"workload": ("configs[2]: SIREN render 800x800, 64+128, ray-sharded" if kind == "siren" else
                                  "render arm of the headline metric: NeRF 800x800, 64+64"),
                     "shard": f"1/{shard} of the image per GPU"}})


def cls_train():
    torch.manual_seed(0)
    nb.set_precision("auto")
    models = [nb.NeRF_3D(CLS=6).to(dev), nb.NeRF_3D(CLS=6).to(dev)]
    emb = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
    n = 8192
    trainer = nb.dp.DataParallelStep(models, emb, 64, 64, False, 1.0, 1.0, 32768, False, cuda_graph=True)
    rays = synth.llff_rays(n, seed=10 + rank).to(dev)
    rgbs = synth.target_rgbs(n, seed=20 + rank).to(dev)
    ms = timed(lambda: trainer.step(rays, rgbs), 5)
    desc = models[0]._desc(False, 10, 4, True)
    arm = "bf16 tcgen05" if nb.ops.mlp_supported(desc, nb._lib.PREC_BF16, True) else "fp32 CUDA cores"
    flops = n * 192 * 4_293_888
    emit({"metric": "rays/s, NeRF_3D (CLS=6) coarse+fine training step (64+64 samples)", "value": n * world / ms * 1e3,
          "unit": "rays/s", "n_gpus": world, "rays_per_gpu": n, "ms_per_step": ms, "mlp_arm": arm,
          "mlp_tflops": flops / ms / 1e9, "config": {"workload": "configs[3]: nerf_cls training, LLFF-shaped rays, batch 8192 / GPU"}})


def siren_train():
    """Not a BASELINE configuration: the SIREN variant through the same training step (kernel (1) backward with sine activations)."""
    torch.manual_seed(0)
    nb.set_precision("auto")
    models = [nb.SirenField().to(dev), nb.SirenField().to(dev)]
    emb = [nb.Embedding(3, 0), nb.Embedding(3, 0)]
    n = 4096
    trainer = nb.dp.DataParallelStep(models, emb, 64, 64, False, 1.0, 1.0, 32768, True, cuda_graph=True)
    rays = synth.blender_rays(n, seed=10 + rank).to(dev)
    rgbs = synth.target_rgbs(n, seed=20 + rank).to(dev)
    ms = timed(lambda: trainer.step(rays, rgbs), 5)
    arm = "bf16 tcgen05" if nb.ops.mlp_supported(models[0]._desc(), nb._lib.PREC_BF16, True) else "fp32 CUDA cores"
    flops = n * 192 * 3 * 2 * 526_848
    emit({"metric": "rays/s, SIREN coarse+fine training step (64+64 samples)", "value": n * world / ms * 1e3, "unit": "rays/s",
          "n_gpus": world, "rays_per_gpu": n, "ms_per_step": ms, "mlp_arm": arm, "mlp_tflops": flops / ms / 1e9,
          "config": {"workload": "SIREN variant of configs[1]: 4096 rays / GPU, 64+64 samples"}})


def render_e2e():
    """The render arm end to end through the public eval-side API (rows f2 + path + f3): host camera pose in, uint8 image and
    PFM depth payload on the host out -- rays generated on the device, 20 chunks of 32k rays, image finished on the device."""
    from nerf_siren_b200 import evaluation as ev
    torch.manual_seed(0)
    nb.set_precision("auto")
    models = [nb.NeRF().to(dev), nb.NeRF().to(dev)]
    emb = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
    c2w = torch.tensor([[1.0, 0.0, 0.0, 0.0], [0.0, 0.0, -1.0, -4.0], [0.0, 1.0, 0.0, 0.0]])
    H = W = 800
    focal = 0.5 * W / 0.3604                                   # camera_angle_x = 0.6911 (lego)
    gt = torch.rand(H * W, 3, device=dev)

    def run():
        r = ev.render_image(models, emb, c2w, H, W, focal, 2.0, 6.0, 64, 64, white_back=True, gt=gt, device=dev)
        return r["img"].cpu(), r["depth_pfm"].cpu(), r["psnr"]
    ms = timed(run, 3)
    emit({"metric": "rays/s, NeRF 800x800 image end to end (pose in, uint8 image + PFM depth on the host out)",
          "value": H * W / ms * 1e3, "unit": "rays/s", "n_gpus": 1, "ms_per_image": ms, "d2h_bytes_per_image": H * W * 7,
          "config": {"workload": "eval.py:152-193 iteration: generate rays, 64+64 render (test_time), quantise / flip / PSNR"}})


for what in (sys.argv[1:] or ["render", "siren", "cls", "siren_train", "render_e2e"]):
    if what == "render_e2e":
        if rank == 0:
            render_e2e()
        continue
    if what in ("render", "siren"):
        render(what)
    elif what == "cls":
        cls_train()
    elif what == "siren_train":
        siren_train()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
