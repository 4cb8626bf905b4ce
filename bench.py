#!/usr/bin/env python
"""bench.py -- the reference's headline metric on B200: rays/s of the NeRF coarse+fine TRAINING step
(BASELINE.json configs[1]: batch 4096 rays per GPU, 64 + 64 samples, bf16 tensor-core MLP, 1/2/4/8 GPUs).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (one rank per GPU under torchrun for N>1)
    python bench.py --impl reference [--steps K] [--warmup W]      # the reference algorithm on the host cores

One "step" = render_rays(coarse 64 + fine 64+64, perturb=1, noise_std=1) -> MSELoss(rgb_coarse)+MSELoss(rgb_fine) ->
backward -> ONE flat gradient all-reduce (N>1) -> Adam, on one batch of synthetic Blender-shaped rays with
random-init NeRF weights.  Prints ONE JSON line (rank 0).  `value`: inputs resident in HBM; `e2e`: the same step
through the public API with HOST (pinned) rays/rgbs copied in and the loss read back every step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_RAYS, N_SAMPLES, N_IMPORTANCE = 4096, 64, 64
MACS_FULL, MACS_SIGMA = 593_408, 491_264                      # MACs per sample, NeRF D=8 W=256
FLOP_FWD_PER_RAY = (N_SAMPLES + N_SAMPLES + N_IMPORTANCE) * 2 * MACS_FULL           # 227.9 MFLOP
FLOP_TRAIN_PER_SAMPLE = 3_489_024                              # fwd + wgrad + dgrad (no dgrad into encodings)
FLOP_TRAIN_PER_RAY = (N_SAMPLES + N_SAMPLES + N_IMPORTANCE) * FLOP_TRAIN_PER_SAMPLE  # 669.9 MFLOP


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


# ------------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, nme in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nme)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        # under load = the upper half of the samples (idle samples before/after the region drag the median down)
        sm.sort()
        return {"sm_mhz": statistics.median(sm[len(sm) // 2:]), "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------
# the reference algorithm on the host cores (oracle port of the reference's torch code; CPU, autograd)
# ------------------------------------------------------------------------------------------------------
def cpu_train_step_factory(n_rays: int):
    """Returns (step_fn, cores): one CPU training step of the reference algorithm on n_rays rays."""
    import torch
    from oracle import render_oracle as O       # checker-side code: used here only as the timed CPU baseline
    torch.set_num_threads(os.cpu_count() or 1)
    pc = {k: v.clone().requires_grad_(True) for k, v in O.nerf_init(seed=1).items()}
    pf = {k: v.clone().requires_grad_(True) for k, v in O.nerf_init(seed=2).items()}
    opt = torch.optim.Adam(list(pc.values()) + list(pf.values()), lr=5e-4, eps=1e-8)
    rays = O.blender_rays(n_rays, seed=3)
    target = torch.rand(n_rays, 3, generator=torch.Generator().manual_seed(4))
    fields = [lambda x, so: O.nerf_mlp(pc, x, so), lambda x, so: O.nerf_mlp(pf, x, so)]

    def step():
        rng = {"perturb_rand": torch.rand(n_rays, N_SAMPLES), "noise_coarse": torch.randn(n_rays, N_SAMPLES),
               "u": torch.rand(n_rays, N_IMPORTANCE), "noise_fine": torch.randn(n_rays, N_SAMPLES + N_IMPORTANCE)}
        opt.zero_grad(set_to_none=True)
        out = O.render_rays(fields, rays, N_SAMPLES, False, 1.0, 1.0, N_IMPORTANCE, 32768, True, False, rng=rng)
        loss = O.mse_loss(out, target)
        loss.backward()
        opt.step()
        return float(loss.detach())
    return step, torch.get_num_threads()


def time_cpu(n_rays: int, steps: int, warmup: int):
    step, cores = cpu_train_step_factory(n_rays)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return n_rays / dt, dt * 1e3, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sample = args.cpu_rays
    value, ms, cores = time_cpu(sample, args.steps, max(1, args.warmup))
    line = {"impl": "reference", "metric": "rays/s, NeRF coarse+fine training step (64+64 samples)", "value": value,
            "unit": "rays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1]: NeRF coarse+fine training step, 4096 rays/GPU, 64+64 samples, bf16",
                       "rays_per_step": sample, "note": "reference algorithm (torch CPU port, oracle/) on the host cores; "
                       "each step is a bounded sample of the workload"},
            "cpu_baseline": {"value": value, "unit": "rays/s", "cores": cores, "kind": "port",
                             "sample": f"{sample} rays x (64 + 128) samples, fwd+bwd+Adam, per step"},
            "e2e": {"value": value, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
def time_kernel(fn, iters=10, warm=3):
    import torch
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def kernel_rooflines(nb, trainer, rays, peaks):
    """Times each hot kernel ALONE on this step's shapes (CUDA events on the launching stream) and relates its
    algorithmic work to the measured peak.  Returns a list of roofline dicts."""
    import torch
    from nerf_siren_b200 import _field, _lib, ops
    dev = rays.device
    n, S, Sf = rays.shape[0], N_SAMPLES, N_SAMPLES + N_IMPORTANCE
    out = []
    mc, mf = trainer.models
    emb_xyz, emb_dir = trainer.embeddings
    desc = mf._desc(False, emb_xyz.N_freqs, emb_dir.N_freqs, emb_xyz.logscale)
    with torch.no_grad():
        flat = _field.flat_params(mf).detach()
        pr = torch.rand(n, S, device=dev)
        z = ops.sample_coarse(rays, S, False, 1.0, pr)
        # (2) sampling: 32 B/ray in + 4*S B/ray rand in + 4*S B/ray out
        ms = time_kernel(lambda: ops.sample_coarse(rays, S, False, 1.0, pr), 50)
        by = n * (32 + 8 * S)
        out.append({"kernel": "sample_coarse_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "ms": ms, "bytes": by})
        prec = _lib.PREC_BF16 if ops.mlp_supported(desc, _lib.PREC_BF16, False) else _lib.PREC_FP32
        ws = ops.mlp_workspace(desc, prec, n * S, False, False, dev)
        raw = ops.mlp_fwd(desc, prec, _lib.IN_RAYS, flat, None, rays, z, n, S, False, False, ws).view(n, S, 4)
        noise = torch.randn(n, S, device=dev)
        w = ops.composite_fwd(raw, z, rays, noise, 1.0, True)[0]
        ms = time_kernel(lambda: ops.composite_fwd(raw, z, rays, noise, 1.0, True), 50)
        by = n * S * (16 + 4 + 4 + 4) + n * (12 + 20)             # raw 16 + z 4 + noise 4 in, w 4 out per sample
        out.append({"kernel": "composite_fwd_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "ms": ms, "bytes": by})
        d_rgb = torch.randn(n, 3, device=dev)
        ms = time_kernel(lambda: ops.composite_bwd(raw, z, rays, noise, 1.0, True, d_rgb=d_rgb), 50)
        by = n * S * (16 + 4 + 4 + 16) + n * (12 + 12)
        out.append({"kernel": "composite_bwd_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "ms": ms, "bytes": by})
        u = torch.rand(n, N_IMPORTANCE, device=dev)
        zf = ops.resample(z, w, N_IMPORTANCE, False, u)
        ms = time_kernel(lambda: ops.resample(z, w, N_IMPORTANCE, False, u), 50)
        by = n * (4 * S + 4 * S + 4 * N_IMPORTANCE + 4 * Sf)
        out.append({"kernel": "resample_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "ms": ms, "bytes": by})
        # (1) MLP forward, fine pass shape (n * 128 points), the arm the step uses
        ws = ops.mlp_workspace(desc, prec, n * Sf, False, False, dev)
        ms = time_kernel(lambda: ops.mlp_fwd(desc, prec, _lib.IN_RAYS, flat, None, rays, zf, n, Sf, False, False, ws), 10)
        fl = n * Sf * 2 * MACS_FULL
        out.append({"kernel": "mlp_tc_fwd_kernel" if prec == _lib.PREC_BF16 else "sgemm_strided_kernel (fp32 fwd chain)",
                    "bound": "tensor", "achieved": fl / ms / 1e9, "peak": peaks["bf16_burst"], "unit": "TFLOP/s", "ms": ms,
                    "flops": fl})
        # MLP forward (saving activations) + backward, fine pass shape, the arm training uses
        tprec = _lib.PREC_BF16 if ops.mlp_supported(desc, _lib.PREC_BF16, True) else _lib.PREC_FP32
        ws = ops.mlp_workspace(desc, tprec, n * Sf, False, True, dev)
        rawf = ops.mlp_fwd(desc, tprec, _lib.IN_RAYS, flat, None, rays, zf, n, Sf, False, True, ws)
        d_raw = torch.randn_like(rawf) * 1e-3
        ms = time_kernel(lambda: ops.mlp_bwd(desc, tprec, _lib.IN_RAYS, flat, None, rays, zf, n, Sf, rawf, d_raw, ws), 5, 2)
        fl = n * Sf * (FLOP_TRAIN_PER_SAMPLE - 2 * MACS_FULL)
        out.append({"kernel": "mlp_tc_bwd (dgrad chain + wgrad)" if tprec == _lib.PREC_BF16 else
                    "sgemm_strided_kernel (fp32 bwd chain)", "bound": "tensor", "achieved": fl / ms / 1e9,
                    "peak": peaks["bf16_burst"], "unit": "TFLOP/s", "ms": ms, "flops": fl})
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp))
    for r in out:
        r["frac"] = r["achieved"] / r["peak"]
        r["traffic"] = traffic.get(r["kernel"])            # DRAM bytes of one launch from the committed ncu capture
        r["peak_source"] = peaks["source"] + (" burst (kernel timed alone)" if r["bound"] == "tensor" else " copy")
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference for the "
                         "CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import nerf_siren_b200 as nb
    from nerf_siren_b200 import synth
    nb.set_precision("auto")
    peaks = load_peaks()

    torch.manual_seed(0)                                   # identical init on every rank (DDP broadcast equivalent)
    models = [nb.NeRF().to(dev), nb.NeRF().to(dev)]
    embeddings = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
    trainer = nb.dp.DataParallelStep(models, embeddings, N_SAMPLES, N_IMPORTANCE, False, 1.0, 1.0, 32768, True,
                                      cuda_graph=not args.no_graph)

    # synthetic data: a pool of distinct batches per rank (each rank draws its own rays, as DistributedSampler would)
    pool = 8
    host_rays = [synth.blender_rays(N_RAYS, seed=1000 * rank + b).pin_memory() for b in range(pool)]
    host_rgbs = [synth.target_rgbs(N_RAYS, seed=1000 * rank + b).pin_memory() for b in range(pool)]
    dev_rays = [r.to(dev) for r in host_rays]
    dev_rgbs = [r.to(dev) for r in host_rgbs]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def step_resident(i):
        trainer.step(dev_rays[i % pool], dev_rgbs[i % pool])

    losses = []

    def step_e2e(i):
        if trainer.cuda_graph:                     # the step copies the pinned host batch straight into its input buffers
            rays, rgbs = host_rays[i % pool], host_rgbs[i % pool]
        else:
            rays = host_rays[i % pool].to(dev, non_blocking=True)
            rgbs = host_rgbs[i % pool].to(dev, non_blocking=True)
        losses.append(trainer.step(rays, rgbs).item())                      # D2H read of the step's loss

    for i in range(args.warmup):
        step_resident(i)
    launches0 = nb.ops.launch_count() + trainer.graph_launches
    with Clocks(local) as clk:
        total_ms = timed(step_resident, args.steps)
    launches = nb.ops.launch_count() + trainer.graph_launches - launches0      # host launches + kernels replayed from the captured step
    for i in range(min(args.warmup, 3)):
        step_e2e(i)
    e2e_ms = timed(step_e2e, args.steps)
    ms_per_step = total_ms / args.steps
    value = N_RAYS * world / (ms_per_step * 1e-3)
    e2e_value = N_RAYS * world / (e2e_ms / args.steps * 1e-3)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    kernels = kernel_rooflines(nb, trainer, dev_rays[0], peaks)
    tensor_k = [k for k in kernels if k["bound"] == "tensor"]
    dominant = max(tensor_k, key=lambda k: k["ms"])
    tensor_ms = sum(k["ms"] for k in tensor_k)
    # the step runs coarse (64 samples) + fine (128 samples): MLP kernels were timed on the fine shape (2/3 of the rows)
    step_tensor_frac = FLOP_TRAIN_PER_RAY * N_RAYS / (ms_per_step * 1e-3) / 1e12 / peaks["bf16_sustained"]
    cpu = None
    if not args.no_cpu_baseline and world == 1:           # the CPU baseline is reported at N = 1 only
        v, ms, cores = time_cpu(args.cpu_rays, 3, 1)
        cpu = {"value": v, "unit": "rays/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_rays} rays x (64 + 128) samples, fwd+bwd+Adam, 3 timed steps after 1 warm-up "
                         f"({ms:.0f} ms/step)"}
    bf16_train = "bf16" in dominant["kernel"] or "tc" in dominant["kernel"]
    line = {
        "metric": "rays/s, NeRF coarse+fine training step (64+64 samples)", "value": value, "unit": "rays/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if all("tc" in k["kernel"] for k in tensor_k) else "bf16 fwd / f32 bwd",
        "data": "synthetic",
        "config": {"workload": "configs[1]: NeRF coarse+fine training step, 4096 rays/GPU, 64+64 samples, bf16",
                   "rays_per_gpu": N_RAYS, "global_rays": N_RAYS * world, "samples": "64 coarse + 128 fine",
                   "parallelism": f"dp{world} (ray-sharded, one flat grad all-reduce per step)",
                   "l2": "per-step working set (saved activations, >1 GB) exceeds the 126 MB L2; 8 distinct batches "
                         "cycled; no explicit flush",
                   "launch": ("forward + backward replayed from one captured CUDA graph per step; all-reduce and Adam launched "
                              "from the host" if trainer.cuda_graph else "every kernel launched from the host")},
        "clocks": clk.summary(),
        "e2e": {"value": e2e_value, "unit": "rays/s", "h2d_bytes_per_step": N_RAYS * 11 * 4 * 1, "d2h_bytes_per_step": 4,
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": launches,
        "roofline": {k: dominant[k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic", "kernel", "ms",
                                              "peak_source")},
        "step_fraction_of_tensor_roofline": step_tensor_frac,
        "kernels": kernels,
        "cpu_baseline": cpu,
        "final_loss": losses[-1] if losses else None,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-rays", type=int, default=256, help="rays per CPU-baseline step (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel of the step from the host instead of replaying the captured forward+backward")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        # torchrun exports OMP_NUM_THREADS=1; the reference arm is meant to use every host core (set before torch loads)
        for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ[k] = str(os.cpu_count() or 1)
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU (the driver launches torchrun itself)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)]
        cmd += sys.argv[1:]
        return subprocess.call(cmd)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
