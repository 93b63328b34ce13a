#!/usr/bin/env python
"""bench.py — DDPM hot-path benchmark (BASELINE.json: train img/s and 1000-step sampled img/s on synthetic
1x28x28 MNIST-shaped data).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [...]                          # the reference algorithm on the host CPU cores
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...   # N>1, one rank per GPU

A "step" is one training step (label drop, eps/u draw, q_sample, UNet fwd, loss, UNet bwd, gradient allreduce,
Adam) over one batch of `--batch` images per GPU (weak scaling).  `value` = images/s with the batch resident in HBM;
`e2e` = the same through the public `train_step` with pinned host batches copied in and the loss read back every
step.  The sampling leg (`sampling`) times the fused ancestral loop (`--sample-steps` steps, unguided) for
`--sample-n` images per GPU, sharded with no communication.  One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "ddpm_train_images_per_sec"
UNIT = "img/s"
FWD_FLOP_PER_IMG = 4.3902e9
TRAIN_FLOP_PER_IMG = 13.169e9


def env_int(k, d):
    return int(os.environ.get(k, d))


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 8] or [r for _, r in self.rows if len(r) >= 8]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[1]) for r in rows)
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": sorted(reasons),
                "power_w_max": max(float(r[3]) for r in rows), "samples": len(rows)}


def synthetic_batch(B, seed, size=28):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 1, size, size, generator=g) * 2 - 1
    y = torch.randint(0, 10, (B,), generator=g)
    return x, y


# ----------------------------------------------------------------------------------------------- reference arm (CPU)
def cpu_reference_rate(sample_B, steps, warmup, threads):
    """The reference algorithm (oracle/ddpm_oracle.py, pinned to the reference's own outputs) on the host cores:
    one train step = loss fwd + autograd bwd + torch.optim.Adam, like diffusion_model.py:63-74 on CPU."""
    from oracle import ddpm_oracle as O
    torch.set_num_threads(threads)
    sd = {k: v.clone().requires_grad_(True) for k, v in O.random_state_dict(1).items()}
    params = list(sd.values())
    opt = torch.optim.Adam(params, lr=3e-4)
    times = []
    for it in range(warmup + steps):
        x, y = synthetic_batch(sample_B, it)
        t0 = time.perf_counter()
        opt.zero_grad()
        eps = torch.randn(x.shape)
        u = torch.rand(sample_B)
        loss, _, _ = O.training_loss(sd, x, y, eps, u, "eps")
        loss.backward()
        opt.step()
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    return sample_B * len(times) / sum(times), sum(times) / len(times)


def cpu_reference_sampling_rate(n, steps, threads):
    from oracle import ddpm_oracle as O
    torch.set_num_threads(threads)
    sd = O.random_state_dict(1)
    y = torch.arange(n) % 10
    init = torch.randn(n, 1, 28, 28)
    noises = [torch.randn(n, 1, 28, 28) for _ in range(steps)]

    def net(z, l, uncond):
        with torch.no_grad():
            return O.unet_forward(sd, z, l, guide=y)

    t0 = time.perf_counter()
    O.sample_loop(net, init, steps, sampler="noisy", mean_type="eps", cond_w=None, noises=noises)
    dt = time.perf_counter() - t0
    return dt / steps  # seconds per sampler step for n images


def run_reference(args, rank):
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    sample_B = args.cpu_batch
    rate, s_per_step = cpu_reference_rate(sample_B, max(1, args.steps), max(1, min(args.warmup, 1)), threads)
    sps = cpu_reference_sampling_rate(4, 3, threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": s_per_step * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"DDPM UNet C=128 1x28x28, train step, batch {args.batch}/GPU fp32/TF32 (configs[1])",
                   "mean_type": "eps", "sampler": "noisy", "timesteps": args.sample_steps},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{args.steps} train steps (fwd+bwd+Adam) of {sample_B} images, torch CPU fp32, "
                                   f"oracle/ddpm_oracle.py (restatement pinned to reference-generated fixtures)"},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "sampling": {"value": 4.0 / (sps * args.sample_steps), "unit": "img/s", "n": 4, "timesteps": args.sample_steps,
                     "extrapolated_from_steps": 3, "guided": False},
    }
    print(json.dumps(line), flush=True)
    return 0
