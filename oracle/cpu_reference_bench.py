#!/usr/bin/env python
"""Time the UNMODIFIED reference on the host CPU cores (BASELINE.json configs[0]) — TEST / BENCH INFRASTRUCTURE ONLY.

    CUDA_VISIBLE_DEVICES="" python oracle/cpu_reference_bench.py --steps 20 --warmup 5 --bs 64 --sample-n 16 --sample-steps 50

What runs is `gms.main`'s own construction path (`load_model_and_data()`, gms/main.py:43-92) with
`--model diffusion_model --device cpu --eval_heavy 0 --mean_type eps --sampler noisy --timesteps 1000 --bs 64
--hidden_size 128` over the synthetic loader, then the inner loop of `gms.main.train` (main.py:205-215: batch
`.to(device)`, `model.train_step`, `.detach().cpu()` of every metric) for `--warmup` + `--steps` batches, timed with
time.perf_counter.  The sampling leg times `GaussianDiffusion.sample` (gaussian_diffusion.py:245-296) of the same model
for `--sample-steps` ancestral steps of `--sample-n` images, unguided (the `evaluate()` path, one net eval per step) and
guided (the `sample(n, y)` path, two), and extrapolates linearly to 1000 steps (every step costs the same).
CUDA is hidden from this process (bench.py launches it with CUDA_VISIBLE_DEVICES=""), so torch.cuda.amp.autocast /
GradScaler are inert exactly as on a CPU-only machine.  Prints one JSON object.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def port_main(a, threads):
    """Fallback when the vendored reference is not on this machine: the same measurements on the oracle port — op for
    op the same torch calls (oracle/ddpm_oracle.py, pinned to reference-generated fixtures), autograd + torch.optim.Adam."""
    import torch
    from oracle import ddpm_oracle as O

    sd = {k: v.clone().requires_grad_(True) for k, v in O.random_state_dict(1).items()}
    opt = torch.optim.Adam(list(sd.values()), lr=3e-4)
    times, last = [], 0.0
    for k in range(a.warmup + a.steps):
        g = torch.Generator().manual_seed(k)
        x = torch.rand(a.bs, 1, 28, 28, generator=g) * 2 - 1
        y = torch.randint(0, 10, (a.bs,), generator=g)
        t0 = time.perf_counter()
        opt.zero_grad()
        y[torch.rand(a.bs) < 0.1] = -1
        loss, _, _ = O.training_loss(sd, x, y, torch.randn(x.shape), torch.rand(a.bs), "eps")
        loss.backward()
        opt.step()
        last = float(loss.detach())
        if k >= a.warmup:
            times.append(time.perf_counter() - t0)
    train = {"img_s": a.bs * len(times) / sum(times), "s_per_step": sum(times) / len(times), "batch": a.bs,
             "steps": len(times), "last_loss": last}
    samp = None
    if a.sample_steps > 0 and a.sample_n > 0:
        n, S = a.sample_n, a.sample_steps
        sdn = {k: v.detach() for k, v in sd.items()}
        y = torch.arange(n) % 10
        init = torch.randn(n, 1, 28, 28)
        noises = [torch.randn(n, 1, 28, 28) for _ in range(S)]
        res = {}
        with torch.no_grad():
            net = lambda z, l, uncond: O.unet_forward(sdn, z, l, guide=(-torch.ones_like(y) if uncond else y))
            for tag, cw in (("unguided", None), ("guided", 4.0 * torch.rand(n))):
                t0 = time.perf_counter()
                O.sample_loop(net, init, S, sampler="noisy", mean_type="eps", cond_w=cw, noises=noises)
                dt = time.perf_counter() - t0
                res[tag] = {"s_per_step": dt / S, "img_s_1000_steps": n / (dt / S * a.timesteps)}
        samp = {"n": n, "measured_steps": S, "extrapolated_to_steps": a.timesteps, **res}
    print(json.dumps({"train": train, "sampling": samp, "threads": threads, "cpu": "", "torch": torch.__version__,
                      "cuda_visible": torch.cuda.is_available(), "reference_root": None, "kind": "port"}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--bs", type=int, default=64)
    ap.add_argument("--sample-n", type=int, default=16)
    ap.add_argument("--sample-steps", type=int, default=50)
    ap.add_argument("--timesteps", type=int, default=1000)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--port", action="store_true",
                    help="time oracle/ddpm_oracle.py (the pinned restatement) instead of the reference; chosen automatically "
                         "when baseline/_ref is absent")
    a = ap.parse_args()

    import warnings
    warnings.filterwarnings("ignore")
    import torch
    from oracle import reference_runtime as R

    threads = a.threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    if a.port or not R.available():
        return port_main(a, threads)
    R.install(synthetic_mnist=True, n_train=max(a.bs * (a.steps + a.warmup + 1), 1024), n_test=max(a.bs, 64))
    tmp = tempfile.mkdtemp(prefix="gms_cpu_ref_")
    subprocess.run(["git", "init", "-q", tmp], check=False)  # common.dump_logger wants a git work tree (unused here)
    os.chdir(tmp)
    import gms.main as gms_main

    sys.argv = ["gms.main", "--model", "diffusion_model", "--device", "cpu", "--eval_heavy", "0", "--mean_type", "eps",
                "--sampler", "noisy", "--timesteps", str(a.timesteps), "--bs", str(a.bs), "--hidden_size", "128",
                "--logdir", os.path.join(tmp, "logs")]
    torch.manual_seed(0)
    model, train_ds, test_ds, _, _, G = gms_main.load_model_and_data()
    assert type(model).__module__ == "gms.diffusion.diffusion_model", type(model)

    # ---- training leg: the inner loop of gms.main.train (main.py:205-215)
    model.train()
    times = []
    it = iter(train_ds)
    last = None
    for k in range(a.warmup + a.steps):
        batch = next(it)
        t0 = time.perf_counter()
        train_x, train_y = batch[0].to(G.device), batch[1].to(G.device)
        metrics = model.train_step(train_x, train_y)
        last = {key: metrics[key].detach().cpu() for key in metrics}
        dt = time.perf_counter() - t0
        if k >= a.warmup:
            times.append(dt)
    train = {"img_s": a.bs * len(times) / sum(times), "s_per_step": sum(times) / len(times), "batch": a.bs,
             "steps": len(times), "last_loss": float(last["loss"])}

    # ---- sampling leg: GaussianDiffusion.sample of the same net for a bounded number of steps, extrapolated
    samp = None
    if a.sample_steps > 0 and a.sample_n > 0:
        from functools import partial

        from gms.diffusion.gaussian_diffusion import GaussianDiffusion

        model.eval()
        n, S = a.sample_n, a.sample_steps
        d = GaussianDiffusion(mean_type="eps", num_steps=S, sampler="noisy", sample_cond_w=-1.0)
        y = torch.arange(n) % 10
        init = torch.randn(n, 1, 28, 28)
        res = {}
        with torch.no_grad():
            for tag, cw in (("unguided", None), ("guided", 0.5)):
                t0 = time.perf_counter()
                d.sample(net=partial(model.net, guide=y), init_x=init, cond_w=cw)
                dt = time.perf_counter() - t0
                res[tag] = {"s_per_step": dt / S, "img_s_1000_steps": n / (dt / S * a.timesteps)}
        samp = {"n": n, "measured_steps": S, "extrapolated_to_steps": a.timesteps, **res}

    cpu = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                cpu = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    print(json.dumps({"train": train, "sampling": samp, "threads": threads, "cpu": cpu, "torch": torch.__version__,
                      "cuda_visible": torch.cuda.is_available(), "reference_root": R.find_root()}))


if __name__ == "__main__":
    main()
