#!/usr/bin/env python
"""bench.py -- the reference's headline metric on B200: "rays/s train & render (64+64 samples) at 1/2/4/8 B200".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config C]      # our arm (one rank per GPU under torchrun for N>1)
    python bench.py --impl reference [--config C] [--steps K] [--warmup W] # the reference algorithm on the host cores

--config (default `train`, the headline = BASELINE.json configs[1]):
    train         NeRF coarse+fine TRAINING step, 4096 rays / GPU, 64+64 samples, bf16 tensor-core MLP      (weak scaling)
                  -- the default line also carries `extra.render`: the NeRF 800x800 64+64 test_time render on this GPU
    render        NeRF 800x800 full-image render, 64+64, test_time=True, rows sharded over the GPUs          (strong)
    siren_render  configs[2]: SIREN variant, 800x800, 64+128, test_time=True, rows sharded over the GPUs     (strong)
    cls_train     configs[3]: NeRF_3D (CLS=6) semantic-head training (render_rays_3d + MSECELoss), LLFF-shaped rays,
                  8192 rays / GPU                                                                            (weak)
    eg3d          configs[4]: EG3D tri-plane renderer, 3x32x256^2 planes, 128x128 rays, 48+48 samples, one replica / GPU
    eg3d_train    configs[4] shape, training step (system.py:111-120): planes + OSGDecoder trained on 16,384 rays / step

One "step" = one pass of the hot path over one batch.  `value`: inputs resident in HBM; `e2e`: the same step through the
public API with HOST (pinned) inputs copied in and the result read back every step.  Every line carries `roofline` (dominant
kernel, timed alone with CUDA events), `cpu_baseline` (N=1 only), `clocks`, `gpu_launches`, and `sustained` (the same step
repeated for >= 2 s, so the power-capped steady state is on record next to the K-step number).
"""
from __future__ import annotations

import argparse
import json
import math
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
MACS_SIREN, MACS_SIREN_SIGMA = 526_848, 459_776               # SIREN full / sigma-only
FLOP_FWD_PER_RAY = (N_SAMPLES + N_SAMPLES + N_IMPORTANCE) * 2 * MACS_FULL           # 227.9 MFLOP
FLOP_TRAIN_PER_SAMPLE = 3_489_024                              # fwd + wgrad + dgrad (no dgrad into encodings)
FLOP_TRAIN_PER_RAY = (N_SAMPLES + N_SAMPLES + N_IMPORTANCE) * FLOP_TRAIN_PER_SAMPLE  # 669.9 MFLOP
# NeRF_3D (CLS = 6) as EXECUTED: parse.1 o parse.0 has no activation in between (nerf_cls.py:56-61), so the kernels run the
# merged 16-column head: 593,408 + 16 * 256 MACs forward (SURVEY 8d's 727,552 is the un-merged count and would overstate).
MACS_CLS = MACS_FULL + 16 * 256
FLOP_CLS_TRAIN_PER_SAMPLE = FLOP_TRAIN_PER_SAMPLE + 3 * 2 * 16 * 256
IMG = 800                                                     # Blender image side of the render configs
EG3D_RAYS, EG3D_S = 128 * 128, 48


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


# ------------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------------
# the inference forward is the tile-pair kernel unless NERFB_FWD_PAIR=0 (csrc/mlp_tc.cu: fwd_pair_mode)
INFER_KERNEL = {"0": "mlp_tc_fwd_kernel", "2": "mlp_tc_fwd_pair2_kernel (cta_group::2)"}.get(os.environ.get("NERFB_FWD_PAIR", "1"),
                                                                                            "mlp_tc_fwd_pair_kernel")
# the backward of the plain NeRF is the fused launch unless NERFB_BWD_FUSED=0 (csrc/mlp_tc_bwd.cu: fused_plan); NeRF_3D / SIREN
# keep dgrad -> wgrad unless NERFB_BWD_FUSED=1
def bwd_kernel_name(cls):
    mode = os.environ.get("NERFB_BWD_FUSED", "2")
    fused = mode in ("1", "2")          # "2" (default): fused for NeRF and NeRF_3D, 3 launches for SIREN (no SIREN training config here)
    return "mlp_tc_bwd_fused_kernel (dgrad chain + wgrad CTAs, one launch)" if fused else "mlp_tc_bwd (dgrad chain + wgrad)"

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


def load_traffic():
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    return json.load(open(tp)) if os.path.exists(tp) else {}


def finish_roofline(r, peaks):
    r["frac"] = r["achieved"] / r["peak"]
    r.setdefault("traffic", load_traffic().get(r["kernel"]))       # DRAM bytes of one launch from the committed ncu capture
    r["peak_source"] = peaks["source"] + (" burst (kernel timed alone)" if r["bound"] == "tensor" else " copy")
    return r


# ------------------------------------------------------------------------------------------------------
# the reference algorithm on the host cores (oracle port of the reference's torch code; CPU).  The ONLY place bench.py
# touches oracle/: as the timed CPU baseline / reference arm, never on the measured GPU path.
# ------------------------------------------------------------------------------------------------------
def _cpu_threads(n=None):
    import torch
    torch.set_num_threads(n or (os.cpu_count() or 1))
    return torch.get_num_threads()


def cpu_step_factory(config: str, n_rays: int):
    """Returns (step_fn, units_per_step, sample_text): one CPU step of the reference algorithm for `config` on a bounded
    sample of the workload."""
    import torch
    from oracle import render_oracle as O
    if config == "train":
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
        return step, n_rays, f"{n_rays} rays x (64 + 128) samples, fwd+bwd+Adam, per step"
    if config in ("cls_train", "cls_train_strong"):
        pc = {k: v.clone().requires_grad_(True) for k, v in O.nerf_init(seed=1, n_cls=6).items()}
        pf = {k: v.clone().requires_grad_(True) for k, v in O.nerf_init(seed=2, n_cls=6).items()}
        opt = torch.optim.Adam(list(pc.values()) + list(pf.values()), lr=5e-4, eps=1e-8)
        rays = O.llff_rays(n_rays, seed=3)
        target = torch.rand(n_rays, 3, generator=torch.Generator().manual_seed(4))
        labels = torch.randint(0, 6, (n_rays,), generator=torch.Generator().manual_seed(5))
        fields = [lambda x, so: O.nerf_mlp(pc, x, so), lambda x, so: O.nerf_mlp(pf, x, so)]

        def step():
            rng = {"perturb_rand": torch.rand(n_rays, N_SAMPLES), "noise_coarse": torch.randn(n_rays, N_SAMPLES),
                   "u": torch.rand(n_rays, N_IMPORTANCE), "noise_fine": torch.randn(n_rays, N_SAMPLES + N_IMPORTANCE)}
            opt.zero_grad(set_to_none=True)
            out = O.render_rays(fields, rays, N_SAMPLES, False, 1.0, 1.0, N_IMPORTANCE, 32768, False, False, rng=rng)
            ce = torch.nn.functional.cross_entropy                       # losses.py:23-59 with weight = 0.5
            loss = 0.5 * O.mse_loss(out, target) + 0.5 * (ce(out["cls_coarse"], labels, ignore_index=-1) +
                                                          ce(out["cls_fine"], labels, ignore_index=-1))
            loss.backward()
            opt.step()
        return step, n_rays, f"{n_rays} LLFF-shaped rays x (64 + 128) samples, NeRF_3D(CLS=6) fwd+bwd+Adam, per step"
    if config == "render":
        pc, pf = O.nerf_init(seed=1), O.nerf_init(seed=2)
        rays = O.blender_rays(n_rays, seed=3)
        fields = [lambda x, so: O.nerf_mlp(pc, x, so), lambda x, so: O.nerf_mlp(pf, x, so)]

        def step():
            with torch.no_grad():
                O.render_rays(fields, rays, 64, False, 0.0, 0.0, 64, 32768, True, True)
        return step, n_rays, f"{n_rays} rays x (64 sigma-only + 128) samples, forward (test_time), per step"
    if config == "siren_render":
        ps = O.siren_init(seed=1)
        rays = O.blender_rays(n_rays, seed=3)
        zero = torch.zeros(1, 9 * 256)

        def field(x, sigma_only):
            y = O.siren_mlp(ps, x[None, :, :3], zero, zero, x[None, :, 3:6] if not sigma_only else x[None, :, :3])[0]
            return y[:, 3:4] if sigma_only else y

        def step():
            with torch.no_grad():
                O.render_rays([field, field], rays, 64, False, 0.0, 0.0, 128, 32768, True, True, n_freq_xyz=0, n_freq_dir=0)
        return step, n_rays, f"{n_rays} rays x (64 sigma-only + 192) samples, SIREN forward (test_time), per step"
    if config == "eg3d":
        from oracle import eg3d_oracle as E
        planes, dec = E.planes_init(seed=5, res=256), E.decoder_init(seed=6)
        ro, rd = E.camera_rays(n_rays, seed=7)
        opts = {"depth_resolution": EG3D_S, "depth_resolution_importance": EG3D_S, "ray_start": 2.2, "ray_end": 3.3,
                "box_warp": 1.0, "white_back": False, "disparity_space_sampling": False, "clamp_mode": "softplus"}
        g = torch.Generator().manual_seed(1)
        rc, u = torch.rand(1, n_rays, EG3D_S, 1, generator=g), torch.rand(n_rays, EG3D_S, generator=g)

        def step():
            with torch.no_grad():
                E.importance_render(planes, dec, ro, rd, opts, rc, u)
        return step, n_rays, f"{n_rays} rays x (48 + 48) samples, tri-plane renderer forward, per step"
    if config == "eg3d_train":
        from oracle import eg3d_oracle as E
        planes = E.planes_init(seed=5, res=256).requires_grad_(True)
        dec = {k: v.requires_grad_(True) for k, v in E.decoder_init(seed=6).items()}
        ro, rd = E.camera_rays(n_rays, seed=7)
        opts = {"depth_resolution": EG3D_S, "depth_resolution_importance": EG3D_S, "ray_start": 2.2, "ray_end": 3.3,
                "box_warp": 1.0, "white_back": False, "disparity_space_sampling": False, "clamp_mode": "softplus"}
        g = torch.Generator().manual_seed(1)
        rc, u = torch.rand(1, n_rays, EG3D_S, 1, generator=g), torch.rand(n_rays, EG3D_S, generator=g)
        target = torch.rand(n_rays, 3, generator=g)
        opt = torch.optim.Adam([planes, *dec.values()], lr=1e-3)

        def step():
            opt.zero_grad()
            out = E.importance_render(planes, dec, ro, rd, opts, rc, u)
            (((out[0][0] - target) ** 2).mean() + ((out[3][0] - target) ** 2).mean()).backward()
            opt.step()
        return step, n_rays, f"{n_rays} rays x (48 + 48) samples, tri-plane renderer forward + autograd backward + Adam, per step"
    raise ValueError(config)


def time_cpu(config: str, n_rays: int, steps: int, warmup: int, threads=None):
    cores = _cpu_threads(threads)
    step, units, sample = cpu_step_factory(config, n_rays)
    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        ts.append(time.perf_counter() - t0)
    dt = sum(ts) / len(ts)
    return {"value": units / dt, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample, "ms_per_step": dt * 1e3,
            "best": units / min(ts), "median": units / statistics.median(ts), "reps": steps}


def cpu_config0():
    """BASELINE.md section 2 / BASELINE.json configs[0]: the reference's render_rays FORWARD on the host cores -- 1024 synthetic
    Blender-shaped rays, NeRF coarse 64 + fine 64, random-init weights, perturb = noise = 0, chunk 32768, test_time=False,
    fp32, no_grad: 1 warm-up + 5 timed repetitions on all threads (best and median), plus a 1-thread figure on a 128-ray
    sample, with the CPU model string."""
    import torch
    from oracle import render_oracle as O
    pc, pf = O.nerf_init(seed=0), O.nerf_init(seed=1)
    fields = [lambda x, so: O.nerf_mlp(pc, x, so), lambda x, so: O.nerf_mlp(pf, x, so)]

    def run(n, reps, threads):
        cores = _cpu_threads(threads)
        rays = O.blender_rays(n, seed=3)
        ts = []
        with torch.no_grad():
            O.render_rays(fields, rays, 64, False, 0.0, 0.0, 64, 32768, True, False)
            for _ in range(reps):
                t0 = time.perf_counter()
                O.render_rays(fields, rays, 64, False, 0.0, 0.0, 64, 32768, True, False)
                ts.append(time.perf_counter() - t0)
        return cores, n / min(ts), n / statistics.median(ts)
    cores, best, med = run(1024, 5, None)
    _, best1, med1 = run(128, 2, 1)
    _cpu_threads(None)
    return {"workload": "configs[0]: render_rays forward on CPU, NeRF coarse64+fine64, 1024 synthetic Blender-shaped rays, "
                        "random-init weights, fp32, no_grad", "unit": "rays/s", "cores": cores, "cpu_model": cpu_model(),
            "best": best, "median": med, "reps": 5, "one_thread": {"rays": 128, "best": best1, "median": med1, "reps": 2},
            "kind": "port"}


METRICS = {
    "train": "rays/s, NeRF coarse+fine training step (64+64 samples)",
    "render": "rays/s, NeRF full-image render 800x800 (64+64 samples, test_time)",
    "siren_render": "rays/s, SIREN full-image render 800x800 (64+128 samples, test_time)",
    "cls_train": "rays/s, NeRF_3D (CLS=6) semantic-head coarse+fine training step (64+64 samples)",
    "eg3d": "rays/s, EG3D tri-plane renderer (3x32x256^2 planes, 48+48 samples)",
    "cls_train_strong": "rays/s, NeRF_3D (CLS=6) semantic-head training step, 8192 rays GLOBAL batch (64+64 samples)",
    "eg3d_train": "rays/s, EG3D tri-plane renderer training step (3x32x256^2 planes + OSGDecoder, 48+48 samples)",
}
WORKLOADS = {
    "train": "configs[1]: NeRF coarse+fine training step, 4096 rays/GPU, 64+64 samples, bf16",
    "render": "render arm of the headline metric: NeRF 800x800 full image, 64+64 samples, test_time, rows sharded over the GPUs",
    "siren_render": "configs[2]: SIREN-MLP variant full-image render 800x800, 64+128 samples, ray-sharded over the GPUs",
    "cls_train": "configs[3]: nerf_cls semantic-head training, LLFF-shaped 504x378 rays, batch 8192 per GPU",
    "eg3d": "configs[4]: EG3D triplane renderer (volumetric_rendering ray_marcher), 3x32x256^2 planes, 96 samples/ray, 128x128 rays",
    "cls_train_strong": "configs[3], strong scaling: nerf_cls semantic-head training, LLFF-shaped rays, 8192 rays global batch split over the GPUs",
    "eg3d_train": "configs[4] shape, training: planes 3x32x256^2 + OSGDecoder optimised on 128x128 rays, 96 samples/ray, MSE on coarse + fine colours, Adam",
}
CPU_SAMPLE_RAYS = {"train": 256, "cls_train": 256, "cls_train_strong": 256, "render": 1024, "siren_render": 1024, "eg3d": 1024,
                   "eg3d_train": 512}
SCALING = {"train": "weak", "cls_train": "weak", "cls_train_strong": "strong", "render": "strong", "siren_render": "strong",
           "eg3d": "weak", "eg3d_train": "weak"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cfg = args.config
    sample = args.cpu_rays or CPU_SAMPLE_RAYS[cfg]
    r = time_cpu(cfg, sample, args.steps, max(1, args.warmup))
    line = {"impl": "reference", "metric": METRICS[cfg], "value": r["value"], "unit": "rays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": SCALING[cfg], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS[cfg], "rays_per_step": sample,
                       "note": "reference algorithm (torch CPU port, oracle/) on the host cores; each step is a bounded "
                               "sample of the workload (rays/s on the CPU is batch-size independent at these sizes)"},
            "cpu_baseline": {"value": r["value"], "unit": "rays/s", "cores": r["cores"], "kind": "port", "sample": r["sample"],
                             "cpu_model": cpu_model(), "best": r["best"], "median": r["median"]},
            "e2e": {"value": r["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if cfg == "train":
        line["config0"] = cpu_config0()
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------------
# our arm: one Workload per --config (what a step is, its resident and host-buffer forms, its dominant-kernel roofline)
# ------------------------------------------------------------------------------------------------------
class Ctx:
    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference for "
                             "the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        import nerf_siren_b200 as nb
        self.nb = nb
        nb.set_precision("auto")
        self.peaks = load_peaks()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps):
        """EXACTLY `steps` steps bracketed by barrier + synchronize, CUDA events on the launching stream, MAX over ranks."""
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        self.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        return float(ms.item())

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


class TrainWorkload:
    """configs[1] (NeRF) / configs[3] (NeRF_3D semantic head): DataParallelStep on this rank's shard of rays."""

    def __init__(self, cx: Ctx, cls: bool, no_graph: bool, strong: bool = False):
        torch, nb = cx.torch, cx.nb
        from nerf_siren_b200 import synth
        self.cx, self.cls = cx, cls
        self.n = 8192 if cls else N_RAYS
        if strong:                                             # configs[3] also at a FIXED global batch of 8192 rays
            self.n = self.n // cx.world
        torch.manual_seed(0)                                   # identical init on every rank (DDP broadcast equivalent)
        if cls:
            models = [nb.NeRF_3D(CLS=6).to(cx.dev), nb.NeRF_3D(CLS=6).to(cx.dev)]
            kw = dict(render_fn=nb.render_rays_3d, render_kwargs={"no_grad_on_nerf": False}, loss_fn=nb.losses.MSECELoss(),
                      loss_kwargs={"weight": 0.5})
            white = False
        else:
            models = [nb.NeRF().to(cx.dev), nb.NeRF().to(cx.dev)]
            kw, white = {}, True
        self.embeddings = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
        self.trainer = nb.dp.DataParallelStep(models, self.embeddings, N_SAMPLES, N_IMPORTANCE, False, 1.0, 1.0, 32768, white,
                                              cuda_graph=not no_graph, **kw)
        pool = self.pool = 8                                   # distinct batches per rank (each rank draws its own rays)
        gen = synth.llff_rays if cls else synth.blender_rays
        self.h_rays = [gen(self.n, seed=1000 * cx.rank + b).pin_memory() for b in range(pool)]
        self.h_rgbs = [synth.target_rgbs(self.n, seed=1000 * cx.rank + b).pin_memory() for b in range(pool)]
        self.h_lab = [torch.randint(0, 6, (self.n, 1), generator=torch.Generator().manual_seed(b)).pin_memory()
                      for b in range(pool)] if cls else [None] * pool
        self.d_rays = [r.to(cx.dev) for r in self.h_rays]
        self.d_rgbs = [r.to(cx.dev) for r in self.h_rgbs]
        self.d_lab = [l.to(cx.dev) if l is not None else None for l in self.h_lab]
        self.losses = []
        self.units_per_step = self.n * cx.world
        self.h2d, self.d2h = self.n * (11 * 4 + (8 if cls else 0)), 4
        self.flop_per_ray = 192 * (FLOP_CLS_TRAIN_PER_SAMPLE if cls else FLOP_TRAIN_PER_SAMPLE)

    def step_resident(self, i):
        b = i % self.pool
        self.trainer.step(self.d_rays[b], self.d_rgbs[b], self.d_lab[b])

    def step_e2e(self, i):
        b = i % self.pool
        tr, dev = self.trainer, self.cx.dev
        if tr.cuda_graph:                     # the step copies the pinned host batch straight into its input buffers
            loss = tr.step(self.h_rays[b], self.h_rgbs[b], self.h_lab[b])
        else:
            lab = self.h_lab[b].to(dev, non_blocking=True) if self.cls else None
            loss = tr.step(self.h_rays[b].to(dev, non_blocking=True), self.h_rgbs[b].to(dev, non_blocking=True), lab)
        self.losses.append(loss.item())                                      # D2H read of the step's loss

    def launches(self):
        return self.cx.nb.ops.launch_count() + self.trainer.graph_launches

    def config(self):
        w = self.cx.world
        return {"rays_per_gpu": self.n, "global_rays": self.n * w, "samples": "64 coarse + 128 fine",
                "parallelism": f"dp{w} (ray-sharded, one flat grad all-reduce per step)",
                "loss": "MSECELoss(weight=0.5) on rgb + composited NeRF_3D class maps, labels ~ randint(0,6)" if self.cls
                        else "MSELoss(rgb_coarse) + MSELoss(rgb_fine)",
                "l2": "per-step working set (saved activations, > 1 GB) exceeds the 126 MB L2; 8 distinct batches cycled; no "
                      "explicit flush",
                "launch": ("forward + backward replayed from one captured CUDA graph per step; all-reduce and Adam launched "
                           "from the host" if self.trainer.cuda_graph else "every kernel launched from the host")}

    def rooflines(self):
        """Each hot kernel ALONE on this step's shapes (CUDA events), algorithmic work against the measured peak."""
        torch, nb, peaks = self.cx.torch, self.cx.nb, self.cx.peaks
        from nerf_siren_b200 import _field, _lib, ops
        dev, n, S, Sf = self.cx.dev, self.n, N_SAMPLES, N_SAMPLES + N_IMPORTANCE
        rays = self.d_rays[0]
        out = []
        mf = self.trainer.models[1]
        desc = mf._desc(False, 10, 4, True)
        C = 4 + (6 if self.cls else 0)
        macs = MACS_CLS if self.cls else MACS_FULL
        with torch.no_grad():
            flat = _field.flat_params(mf).detach()
            pr = torch.rand(n, S, device=dev)
            z = ops.sample_coarse(rays, S, False, 1.0, pr)
            ms = time_kernel(lambda: ops.sample_coarse(rays, S, False, 1.0, pr), 50)
            by = n * (32 + 8 * S)                                   # 32 B/ray in + 4*S rand in + 4*S out
            out.append({"kernel": "sample_coarse_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "ms": ms, "bytes": by})
            prec = _lib.PREC_BF16 if ops.mlp_supported(desc, _lib.PREC_BF16, False) else _lib.PREC_FP32
            ws = ops.mlp_workspace(desc, prec, n * S, False, False, dev)
            raw = ops.mlp_fwd(desc, prec, _lib.IN_RAYS, flat, None, rays, z, n, S, False, False, ws).view(n, S, C)
            noise = torch.randn(n, S, device=dev)
            w = ops.composite_fwd(raw, z, rays, noise, 1.0, True)[0]
            ms = time_kernel(lambda: ops.composite_fwd(raw, z, rays, noise, 1.0, True), 50)
            by = n * S * (4 * C + 4 + 4 + 4) + n * (12 + 20 + 4 * (C - 4))
            out.append({"kernel": "composite_fwd_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "ms": ms, "bytes": by})
            d_rgb = torch.randn(n, 3, device=dev)
            ms = time_kernel(lambda: ops.composite_bwd(raw, z, rays, noise, 1.0, True, d_rgb=d_rgb), 50)
            by = n * S * (4 * C + 4 + 4 + 4 * C) + n * (12 + 12)
            out.append({"kernel": "composite_bwd_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "ms": ms, "bytes": by})
            u = torch.rand(n, N_IMPORTANCE, device=dev)
            zf = ops.resample(z, w, N_IMPORTANCE, False, u)
            ms = time_kernel(lambda: ops.resample(z, w, N_IMPORTANCE, False, u), 50)
            by = n * (4 * S + 4 * S + 4 * N_IMPORTANCE + 4 * Sf)
            out.append({"kernel": "resample_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "ms": ms, "bytes": by})
            ws = ops.mlp_workspace(desc, prec, n * Sf, False, False, dev)
            ms = time_kernel(lambda: ops.mlp_fwd(desc, prec, _lib.IN_RAYS, flat, None, rays, zf, n, Sf, False, False, ws), 10)
            fl = n * Sf * 2 * macs
            out.append({"kernel": INFER_KERNEL if prec == _lib.PREC_BF16 else "sgemm_strided_kernel (fp32 fwd chain)",
                        "bound": "tensor", "achieved": fl / ms / 1e9, "peak": peaks["bf16_burst"], "unit": "TFLOP/s", "ms": ms,
                        "flops": fl})
            tprec = _lib.PREC_BF16 if ops.mlp_supported(desc, _lib.PREC_BF16, True) else _lib.PREC_FP32
            ws = ops.mlp_workspace(desc, tprec, n * Sf, False, True, dev)
            ms = time_kernel(lambda: ops.mlp_fwd(desc, tprec, _lib.IN_RAYS, flat, None, rays, zf, n, Sf, False, True, ws), 10)
            out.append({"kernel": "mlp_tc_fwd_kernel (saving activations)" if tprec == _lib.PREC_BF16 else
                        "sgemm_strided_kernel (fp32 fwd chain, saving)", "bound": "tensor", "achieved": fl / ms / 1e9,
                        "peak": peaks["bf16_burst"], "unit": "TFLOP/s", "ms": ms, "flops": fl})
            rawf = ops.mlp_fwd(desc, tprec, _lib.IN_RAYS, flat, None, rays, zf, n, Sf, False, True, ws)
            d_raw = torch.randn_like(rawf) * 1e-3
            ms = time_kernel(lambda: ops.mlp_bwd(desc, tprec, _lib.IN_RAYS, flat, None, rays, zf, n, Sf, rawf, d_raw, ws), 5, 2)
            fl = n * Sf * ((FLOP_CLS_TRAIN_PER_SAMPLE if self.cls else FLOP_TRAIN_PER_SAMPLE) - 2 * macs)
            out.append({"kernel": bwd_kernel_name(self.cls) if tprec == _lib.PREC_BF16 else
                        "sgemm_strided_kernel (fp32 bwd chain)", "bound": "tensor", "achieved": fl / ms / 1e9,
                        "peak": peaks["bf16_burst"], "unit": "TFLOP/s", "ms": ms, "flops": fl})
        return [finish_roofline(r, peaks) for r in out]


class RenderWorkload:
    """Full-image render (test_time=True), rows of ONE 800 x 800 view sharded contiguously over the ranks."""

    def __init__(self, cx: Ctx, siren: bool, shard_of: int = 0):
        torch, nb = cx.torch, cx.nb
        from nerf_siren_b200 import synth
        self.cx, self.siren = cx, siren
        torch.manual_seed(0)
        if siren:
            self.models = [nb.SirenField().to(cx.dev), nb.SirenField().to(cx.dev)]
            self.emb = [nb.Embedding(3, 0), nb.Embedding(3, 0)]
            self.S, self.Si, self.mc, self.mf = 64, 128, MACS_SIREN_SIGMA, MACS_SIREN
        else:
            self.models = [nb.NeRF().to(cx.dev), nb.NeRF().to(cx.dev)]
            self.emb = [nb.Embedding(3, 10), nb.Embedding(3, 4)]
            self.S, self.Si, self.mc, self.mf = 64, 64, MACS_SIGMA, MACS_FULL
        rays = synth.blender_image_rays(IMG, IMG, seed=1)
        shards = shard_of or cx.world
        lo, hi = nb.dp.shard_bounds(rays.shape[0], cx.rank if not shard_of else 0, shards)
        self.h_rays = rays[lo:hi].contiguous().pin_memory()
        self.d_rays = self.h_rays.to(cx.dev)
        self.n = hi - lo
        self.units_per_step = rays.shape[0] if not shard_of else self.n
        self.h2d, self.d2h = self.n * 32, self.n * 20
        self.flop_per_ray = 2 * (self.S * self.mc + (self.S + self.Si) * self.mf)
        self.out_host = torch.empty(self.n, 5).pin_memory()
        self.chunk = 32768 * 4

    def _render(self, rays):
        with self.cx.torch.no_grad():
            return self.cx.nb.render_rays(self.models, self.emb, rays, self.S, False, 0, 0, self.Si, 32768, True, test_time=True)

    def step_resident(self, i):
        self._render(self.d_rays)

    def step_e2e(self, i):
        """eval.py:85-103 shape: host rays in chunks -> device -> render -> rgb, depth, opacity back on the host."""
        torch = self.cx.torch
        for lo in range(0, self.n, self.chunk):
            r = self.h_rays[lo:lo + self.chunk].to(self.cx.dev, non_blocking=True)
            o = self._render(r)
            m = r.shape[0]
            self.out_host[lo:lo + m, 0:3].copy_(o["rgb_fine"], non_blocking=True)
            self.out_host[lo:lo + m, 3].copy_(o["depth_fine"], non_blocking=True)
            self.out_host[lo:lo + m, 4].copy_(o["opacity_fine"], non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def launches(self):
        return self.cx.nb.ops.launch_count()

    def config(self):
        return {"rays_per_gpu": self.n, "image": f"{IMG}x{IMG}", "samples": f"{self.S} coarse (sigma only) + {self.S + self.Si} fine",
                "parallelism": f"rows of the image sharded over {self.cx.world} GPU(s), weights replicated, no collective",
                "l2": "inputs of a step (rays, z, raw: > 300 MB per GPU at N=1) exceed the 126 MB L2; no explicit flush"}

    def rooflines(self):
        torch, nb, peaks = self.cx.torch, self.cx.nb, self.cx.peaks
        from nerf_siren_b200 import _field, _lib, ops
        dev = self.cx.dev
        n, Sf = min(self.n, 4096 * 8), self.S + self.Si
        rays = self.d_rays[:n].contiguous()
        mf = self.models[1]
        desc = mf._desc(False, 0, 0, True) if self.siren else mf._desc(False, 10, 4, True)
        film = mf._film() if self.siren else None
        out = []
        with torch.no_grad():
            flat = _field.flat_params(mf).detach()
            zf = torch.sort(torch.rand(n, Sf, device=dev) * 4 + 2, -1)[0].contiguous()
            prec = _lib.PREC_BF16 if ops.mlp_supported(desc, _lib.PREC_BF16, False) else _lib.PREC_FP32
            ws = ops.mlp_workspace(desc, prec, n * Sf, False, False, dev)
            ms = time_kernel(lambda: ops.mlp_fwd(desc, prec, _lib.IN_RAYS, flat, film, rays, zf, n, Sf, False, False, ws), 10)
            fl = n * Sf * 2 * self.mf
            out.append({"kernel": (INFER_KERNEL + "<SIREN>" if self.siren else INFER_KERNEL) if prec == _lib.PREC_BF16
                        else "sgemm_strided_kernel (fp32 fwd chain)", "bound": "tensor", "achieved": fl / ms / 1e9,
                        "peak": peaks["bf16_burst"], "unit": "TFLOP/s", "ms": ms, "flops": fl, "rows": n * Sf})
            w = torch.rand(n, self.S, device=dev)
            z = torch.sort(torch.rand(n, self.S, device=dev) * 4 + 2, -1)[0].contiguous()
            ms = time_kernel(lambda: ops.resample(z, w, self.Si, True, None), 20)
            by = n * (8 * self.S + 4 * Sf)
            out.append({"kernel": "resample_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "ms": ms, "bytes": by})
            raw = torch.rand(n, Sf, 4, device=dev)
            ms = time_kernel(lambda: ops.composite_fwd(raw, zf, rays, None, 0.0, True), 20)
            by = n * Sf * (16 + 4 + 4) + n * 32
            out.append({"kernel": "composite_fwd_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "ms": ms, "bytes": by})
        return [finish_roofline(r, peaks) for r in out]


class Eg3dWorkload:
    """configs[4]: one 128 x 128 view through kernel (6) in chunks of 4096 rays (eval_eg3d.py:65), 48 + 48 samples; one
    independent replica per GPU (nothing to exchange)."""

    def __init__(self, cx: Ctx):
        torch, nb = cx.torch, cx.nb
        from nerf_siren_b200 import synth
        self.cx = cx
        self.planes = synth.eg3d_planes(256, 32, seed=5).to(cx.dev)
        self.cl = nb.ops.triplane_relayout(self.planes)
        self.dec = synth.eg3d_decoder(seed=6).to(cx.dev)
        o, d = synth.eg3d_camera_rays(EG3D_RAYS, seed=7 + cx.rank)
        self.h_o, self.h_d = o.pin_memory(), d.pin_memory()
        self.d_o, self.d_d = o.to(cx.dev), d.to(cx.dev)
        self.n, self.chunk = EG3D_RAYS, 4096
        self.units_per_step = self.n * cx.world
        self.h2d, self.d2h = self.n * 24, self.n * 20
        self.out_host = torch.empty(self.n, 5).pin_memory()
        self.seed = 0
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=cx.dev)

    def _chunk(self, o, d):
        self.seed += 1
        return self.cx.nb.ops.triplane_render(self.cl, self.dec, 1.0, o, d, EG3D_S, EG3D_S, 2.2, 3.3, None, None, False, 1.0,
                                              False, None, None, self.seed)

    def step_resident(self, i):
        self.flush.zero_()                                   # planes (25 MB) would otherwise stay in the 126 MB L2 across steps
        for lo in range(0, self.n, self.chunk):
            self._chunk(self.d_o[lo:lo + self.chunk], self.d_d[lo:lo + self.chunk])

    def step_e2e(self, i):
        dev = self.cx.dev
        self.flush.zero_()
        for lo in range(0, self.n, self.chunk):
            o = self.h_o[lo:lo + self.chunk].to(dev, non_blocking=True)
            d = self.h_d[lo:lo + self.chunk].to(dev, non_blocking=True)
            res = self._chunk(o, d)
            m = o.shape[0]
            self.out_host[lo:lo + m, 0:3].copy_(res[3], non_blocking=True)
            self.out_host[lo:lo + m, 3].copy_(res[4], non_blocking=True)
            self.out_host[lo:lo + m, 4].copy_(res[5], non_blocking=True)
        self.cx.torch.cuda.current_stream().synchronize()

    def launches(self):
        return self.cx.nb.ops.launch_count()

    def config(self):
        return {"rays_per_gpu": self.n, "chunk": self.chunk, "samples": "48 coarse + 48 importance", "planes": "3 x 32 x 256 x 256 fp32",
                "parallelism": f"{self.cx.world} independent replica(s) (replicas only: nothing to exchange)",
                "l2": "256 MB written before every step (planes start cold in L2)"}

    def rooflines(self):
        peaks = self.cx.peaks
        ms = time_kernel(lambda: self._chunk(self.d_o, self.d_d), 20)
        samples = self.n * 2 * EG3D_S
        by = samples * 1536
        r = {"kernel": "triplane_render_kernel", "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"], "unit": "GB/s",
             "ms": ms, "bytes": by, "note": "algorithmic bytes = 1,536 B of bilinear taps gathered per sample (served from L2: the "
             "25 MB of planes are the only compulsory HBM traffic); all 16,384 rays in one launch"}
        return [finish_roofline(r, peaks)]


class Eg3dTrainWorkload:
    """EG3D training step at the configs[4] shape (the reference's EG3DSystem.training_step, system.py:111-120): 16,384 rays
    of one view against target colours -> EG3D_Renderer.render with grad (one fused launch keeping its samples) -> MSE on
    coarse + fine colours -> backward (marcher backward x 2, field backward) -> Adam (kernel 5) on planes and decoder.  One
    independent replica per GPU (the reference trains EG3D on one GPU; nothing to exchange here)."""

    def __init__(self, cx: Ctx):
        torch, nb = cx.torch, cx.nb
        from nerf_siren_b200 import synth
        self.cx = cx
        kw = dict(nb.eg3d.DEFAULT_RENDERING_KWARGS, depth_resolution=EG3D_S, depth_resolution_importance=EG3D_S, ray_start=2.2,
                  ray_end=3.3, box_warp=1.0)
        self.ren = nb.eg3d.EG3D_Renderer(nb.eg3d.LearnedPlanes(resolution=256), kw).to(cx.dev)
        with torch.no_grad():
            self.ren.plane_source.planes.copy_(synth.eg3d_planes(256, 32, seed=5).reshape(1, 96, 256, 256))
        self.params = [p for p in self.ren.parameters()]
        self.m = [torch.zeros_like(p) for p in self.params]
        self.v = [torch.zeros_like(p) for p in self.params]
        o, d = synth.eg3d_camera_rays(EG3D_RAYS, seed=7 + cx.rank)
        g = torch.Generator().manual_seed(11)
        tgt = torch.rand(EG3D_RAYS, 3, generator=g)
        self.h_o, self.h_d, self.h_t = o.pin_memory(), d.pin_memory(), tgt.pin_memory()
        self.d_o, self.d_d, self.d_t = o.to(cx.dev), d.to(cx.dev), tgt.to(cx.dev)
        self.n = EG3D_RAYS
        self.units_per_step = self.n * cx.world
        self.h2d, self.d2h = self.n * 36, 4
        self.t = 0
        self.losses = []
        self.loss_host = torch.empty(1).pin_memory()
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=cx.dev)

    def _step(self, o, d, tgt):
        nb = self.cx.nb
        for p in self.params:
            p.grad = None
        out = self.ren.render(-1, o, d)
        loss = ((out["rgb_coarse"] - tgt) ** 2).mean() + ((out["rgb_fine"] - tgt) ** 2).mean()     # losses.py:8-21 shape
        loss.backward()
        self.t += 1
        for p, m, v in zip(self.params, self.m, self.v):
            nb.ops.adam_step(p.data.view(-1), p.grad.contiguous().view(-1), m.view(-1), v.view(-1), self.t, lr=1e-3)
        return loss.detach()

    def step_resident(self, i):
        self.flush.zero_()
        self._step(self.d_o, self.d_d, self.d_t)

    def step_e2e(self, i):
        dev = self.cx.dev
        self.flush.zero_()
        loss = self._step(self.h_o.to(dev, non_blocking=True), self.h_d.to(dev, non_blocking=True), self.h_t.to(dev, non_blocking=True))
        self.loss_host.copy_(loss.reshape(1), non_blocking=True)
        self.cx.torch.cuda.current_stream().synchronize()
        self.losses.append(float(self.loss_host[0]))

    def launches(self):
        return self.cx.nb.ops.launch_count()

    def config(self):
        return {"rays_per_gpu": self.n, "samples": "48 coarse + 48 importance", "planes": "3 x 32 x 256 x 256 fp32 (trained) + OSGDecoder 2,372 parameters",
                "loss": "MSE(rgb_coarse) + MSE(rgb_fine)", "optimizer": "Adam lr 1e-3 (nerfb_adam_step)",
                "parallelism": f"{self.cx.world} independent replica(s) (replicas only: nothing to exchange)",
                "l2": "256 MB written before every step (planes start cold in L2)"}

    def rooflines(self):
        peaks, nb, torch = self.cx.peaks, self.cx.nb, self.cx.torch
        P = self.n * 2 * EG3D_S
        planes = self.ren.plane_source.planes.detach().view(3, 32, 256, 256)
        cl, dec = nb.ops.triplane_relayout(planes), self.ren.decoder.flat()
        g = torch.Generator(device=self.cx.dev).manual_seed(3)
        with torch.no_grad():
            smp = nb.ops.triplane_render(cl, dec, 1.0, self.d_o, self.d_d, EG3D_S, EG3D_S, 2.2, 3.3, None, None, False, 1.0, False,
                                         None, None, 1, return_samples=True)[6]
        coords = torch.addcmul(self.d_o.unsqueeze(1), smp[:, 0].unsqueeze(-1), self.d_d.unsqueeze(1)).reshape(-1, 3).contiguous()
        d_rgb, d_sig = torch.randn(P, 3, device=self.cx.dev, generator=g), torch.randn(P, device=self.cx.dev, generator=g)
        ms_b = time_kernel(lambda: nb.ops.triplane_field_bwd(cl, dec, 1.0, coords, 1.0, d_rgb, d_sig), 20)
        ms_f = time_kernel(lambda: nb.ops.triplane_render(cl, dec, 1.0, self.d_o, self.d_d, EG3D_S, EG3D_S, 2.2, 3.3, None, None, False,
                                                          1.0, False, None, None, 1, return_samples=True), 20)
        out = []
        for name, ms, by, note in (
                ("triplane_field_bwd_mma_kernel", ms_b, P * 3072, "algorithmic bytes = 1,536 B of bilinear taps gathered + 1,536 B of plane-gradient "
                 "rows reduced per sample (both served by L2; includes the zero-fill of the 25 MB gradient image)"),
                ("triplane_render_kernel", ms_f, P * 1536 + P * 20, "training forward: 1,536 B of taps per sample + 20 B per sample kept for the backward")):
            out.append(finish_roofline({"kernel": name, "bound": "hbm", "achieved": by / ms / 1e6, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                        "ms": ms, "bytes": by, "note": note}, peaks))
        return out


def make_workload(cx: Ctx, config: str, args):
    if config == "train":
        return TrainWorkload(cx, False, args.no_graph)
    if config == "cls_train":
        return TrainWorkload(cx, True, args.no_graph)
    if config == "cls_train_strong":
        return TrainWorkload(cx, True, args.no_graph, strong=True)
    if config == "render":
        return RenderWorkload(cx, False)
    if config == "siren_render":
        return RenderWorkload(cx, True)
    if config == "eg3d":
        return Eg3dWorkload(cx)
    if config == "eg3d_train":
        return Eg3dTrainWorkload(cx)
    raise ValueError(config)


def measure(cx: Ctx, wl, steps: int, warmup: int, sustain_s: float = 2.0):
    """value (resident inputs), e2e (host buffers), sustained (>= sustain_s of back-to-back steps), launches, clocks."""
    for i in range(warmup):
        wl.step_resident(i)
    l0 = wl.launches()
    with Clocks(cx.local) as clk:
        total_ms = cx.timed(wl.step_resident, steps)
    launches = wl.launches() - l0
    ms = total_ms / steps
    for i in range(min(warmup, 3)):
        wl.step_e2e(i)
    e2e_ms = cx.timed(wl.step_e2e, steps) / steps
    sustained = None
    if sustain_s > 0:
        k = max(steps, int(math.ceil(sustain_s * 1e3 / ms)))
        with Clocks(cx.local) as clk2:
            s_ms = cx.timed(wl.step_resident, k) / k
        sustained = {"steps": k, "ms_per_step": s_ms, "value": wl.units_per_step / (s_ms * 1e-3), "unit": "rays/s",
                     "seconds": k * s_ms * 1e-3, "clocks": clk2.summary()}
    return {"ms": ms, "value": wl.units_per_step / (ms * 1e-3), "e2e_ms": e2e_ms, "e2e_value": wl.units_per_step / (e2e_ms * 1e-3),
            "launches": launches, "clocks": clk.summary(), "sustained": sustained}


def run_ours(args):
    cx = Ctx()
    cfg = args.config
    wl = make_workload(cx, cfg, args)
    m = measure(cx, wl, args.steps, args.warmup, args.sustain)
    extra = {}
    if cfg == "train" and not args.no_extra:
        # the render half of "rays/s train & render": this GPU's share of one 800 x 800 NeRF image (64+64, test_time)
        rw = RenderWorkload(cx, False)
        rm = measure(cx, rw, 3, 3, 0.0)
        extra["render"] = {"metric": METRICS["render"], "value": rm["value"], "unit": "rays/s", "ms_per_image_shard": rm["ms"],
                           "rays_per_gpu": rw.n, "n_gpus": cx.world, "scaling": "strong",
                           "e2e": {"value": rm["e2e_value"], "unit": "rays/s", "h2d_bytes_per_step": rw.h2d,
                                   "d2h_bytes_per_step": rw.d2h},
                           "mlp_tflops": rw.flop_per_ray * rw.n / rm["ms"] / 1e9,
                           "frac_of_sustained_bf16_peak": rw.flop_per_ray * rw.n / rm["ms"] / 1e9 / cx.peaks["bf16_sustained"],
                           "gpu_launches": rm["launches"], "steps": 3, "warmup": 3}
    if cx.rank != 0:
        cx.close()
        return 0
    kernels = wl.rooflines()
    tensor_k = [k for k in kernels if k["bound"] == "tensor"]
    dominant = max(tensor_k or kernels, key=lambda k: k["ms"])
    cpu = None
    if not args.no_cpu_baseline and cx.world == 1:           # the CPU baseline is reported at N = 1 only
        r = time_cpu(cfg, args.cpu_rays or CPU_SAMPLE_RAYS[cfg], 3, 1)
        cpu = {"value": r["value"], "unit": "rays/s", "cores": r["cores"], "kind": "port", "cpu_model": cpu_model(),
               "sample": r["sample"] + f", 3 timed steps after 1 warm-up ({r['ms_per_step']:.0f} ms/step)",
               "best": r["best"], "median": r["median"]}
        if cfg == "train":
            cpu["config0"] = cpu_config0()
    bf16 = all(("tc" in k["kernel"]) for k in tensor_k) if tensor_k else False
    line = {
        "metric": METRICS[cfg], "value": m["value"], "unit": "rays/s", "n_gpus": cx.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": m["ms"], "higher_is_better": True, "scaling": SCALING[cfg], "vs_baseline": None,
        "dtype": ("f32" if cfg in ("eg3d", "eg3d_train") else ("bf16" if bf16 else "bf16 fwd / f32 bwd")), "data": "synthetic",
        "config": dict({"workload": WORKLOADS[cfg]}, **wl.config()),
        "clocks": m["clocks"],
        "e2e": {"value": m["e2e_value"], "unit": "rays/s", "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h,
                "ms_per_step": m["e2e_ms"]},
        "gpu_launches": m["launches"],
        "roofline": {k: dominant.get(k) for k in ("bound", "achieved", "peak", "unit", "frac", "traffic", "kernel", "ms",
                                                  "peak_source")},
        "sustained": m["sustained"],
        "kernels": kernels,
        "cpu_baseline": cpu,
    }
    if hasattr(wl, "flop_per_ray"):
        per_gpu = wl.n if hasattr(wl, "n") else 0
        line["step_fraction_of_tensor_roofline"] = wl.flop_per_ray * per_gpu / (m["ms"] * 1e-3) / 1e12 / cx.peaks["bf16_sustained"]
    if getattr(wl, "losses", None):
        line["final_loss"] = wl.losses[-1]
    if extra:
        line["extra"] = extra
    emit(line)
    cx.close()
    return 0


_JSON_OUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints "NCCL version ..." at init on the GPU
    boxes, from C): keep a private handle on the real stdout for the JSON line and point fd 1 at stderr for everything else."""
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
    return _JSON_OUT


def emit(line):
    out = claim_stdout()
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="train", choices=sorted(METRICS))
    ap.add_argument("--cpu-rays", type=int, default=0, help="rays per CPU-baseline step (bounded sample; 0 = per-config default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra.render object of the default line")
    ap.add_argument("--sustain", type=float, default=2.0, help="seconds of back-to-back steps for the `sustained` object (0 = off)")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel of the step from the host instead of replaying the captured forward+backward")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        # torchrun exports OMP_NUM_THREADS=1; the reference arm is meant to use every host core (set before torch loads)
        for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ[k] = str(os.cpu_count() or 1)
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not (args.gpus > 1 and world == 1):
        claim_stdout()                                  # before NCCL / CUDA initialise
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU (the driver launches torchrun itself)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)]
        cmd += sys.argv[1:]
        return subprocess.call(cmd)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
