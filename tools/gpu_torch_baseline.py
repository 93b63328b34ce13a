#!/usr/bin/env python
"""The same-box Blackwell baseline: the UNMODIFIED reference modules in torch eager on one B200 (SURVEY.md §2.3,
BASELINE.md §4) — cuDNN / cuBLAS / ATen kernels, timed with CUDA events exactly like bench.py times this repo.

    python tools/gpu_torch_baseline.py [--batch 512] [--steps 10] [--sample-steps 20] [--profile profiles/x.md]

Variants of the train step (zero_grad, label drop, loss, backward, Adam — diffusion_model.py:63-74):
  ref_amp_fp16   the reference's own `DiffusionModel.train_step` (torch.cuda.amp fp16 autocast + GradScaler)
  tf32           the same step without autocast: fp32 storage, cuDNN TF32 convolutions (torch's default allow_tf32)
  fp32           ... with cudnn.allow_tf32 = False (true fp32: the numerics oracle's precision)
  bf16           torch.autocast(bfloat16), no GradScaler
  *_cl           the same with the net and the input in channels_last memory format
and of one ancestral sampler step (`GaussianDiffusion.reverse_dpm_step`, unguided and guided) under no_grad.
Prints one JSON object; --profile also writes a torch.profiler kernel table of the tf32 train step.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sample-n", type=int, default=512)
    ap.add_argument("--sample-steps", type=int, default=20)
    ap.add_argument("--profile", default="")
    ap.add_argument("--variants", default="ref_amp_fp16,tf32,fp32,bf16,tf32_cl,bf16_cl")
    a = ap.parse_args()
    import warnings
    warnings.filterwarnings("ignore")
    from functools import partial

    from oracle import reference_runtime as R

    dev = "cuda"
    B = a.batch
    g = torch.Generator().manual_seed(0)
    x = (torch.rand(B, 1, 28, 28, generator=g) * 2 - 1).to(dev)
    y = torch.randint(0, 10, (B,), generator=g).to(dev)
    out = {"batch": B, "gpu": torch.cuda.get_device_name(0), "torch": torch.__version__,
           "cudnn": torch.backends.cudnn.version(), "train": {}, "sampler_step": {}}

    def fresh(channels_last):
        torch.manual_seed(0)
        m, G = R.reference_model(dev, mean_type="eps", sampler="noisy", timesteps=1000, bs=B)
        gen = torch.Generator().manual_seed(1)
        for k, p in m.net.named_parameters():  # same non-degenerate fill of the zero-init convs as bench.py
            if ".out_layers.3." in k:
                p.data.copy_((torch.randn(p.shape, generator=gen) * 0.02).to(dev))
        if channels_last:
            m.net.to(memory_format=torch.channels_last)
        return m

    def make_step(m, mode, channels_last):
        xin = x.contiguous(memory_format=torch.channels_last) if channels_last else x
        if mode == "ref_amp_fp16":
            return lambda: m.train_step(xin, y.clone())

        def step():
            m.optimizer.zero_grad()
            yy = y.clone()
            yy[torch.rand(B) < m.G.cf_drop_prob] = -1
            if mode == "bf16":
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    loss, _ = m.loss(xin, yy)
            else:
                loss, _ = m.loss(xin, yy)
            loss.backward()
            m.optimizer.step()
        return step

    for v in a.variants.split(","):
        cl = v.endswith("_cl")
        mode = v[:-3] if cl else v
        torch.backends.cudnn.allow_tf32 = mode != "fp32"
        torch.backends.cuda.matmul.allow_tf32 = False  # torch default; the 16 small Linears do not matter
        m = fresh(cl)
        m.train()
        try:
            ms = timed(make_step(m, mode, cl), a.steps, a.warmup)
            out["train"][v] = {"ms_per_step": ms, "img_s": B / ms * 1e3}
        except Exception as e:  # e.g. out of memory for a variant
            out["train"][v] = {"error": repr(e)[:200]}
        if a.profile and v == "tf32":
            from torch.profiler import ProfilerActivity, profile
            st = make_step(m, mode, cl)
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                st()
                torch.cuda.synchronize()
            os.makedirs(os.path.dirname(os.path.abspath(a.profile)), exist_ok=True)
            with open(a.profile, "w") as f:
                f.write(f"# torch.profiler kernel table: reference train step, torch eager, fp32 storage + cuDNN TF32, "
                        f"batch {B}, {out['gpu']}, torch {torch.__version__}, cuDNN {out['cudnn']}\n\n```\n")
                f.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=90))
                f.write("\n```\n")
        # ---- one ancestral sampler step (gaussian_diffusion.py:215-243) with the same net, no_grad
        if mode in ("tf32", "bf16", "fp32") and a.sample_steps > 0:
            m.eval()
            n = a.sample_n
            z = torch.randn(n, 1, 28, 28, device=dev)
            if cl:
                z = z.contiguous(memory_format=torch.channels_last)
            ys = torch.arange(n, device=dev) % 10
            lt, ls = torch.tensor(0.3, device=dev), torch.tensor(0.31, device=dev)
            for tag, cw in (("unguided", None), ("guided", 4.0 * torch.rand(n, device=dev))):
                def sstep():
                    with torch.no_grad():
                        if mode == "bf16":
                            with torch.autocast("cuda", dtype=torch.bfloat16):
                                m.diffusion.reverse_dpm_step(net=partial(m.net, guide=ys), logsnr_t=lt, logsnr_s=ls, z_t=z, cond_w=cw)
                        else:
                            m.diffusion.reverse_dpm_step(net=partial(m.net, guide=ys), logsnr_t=lt, logsnr_s=ls, z_t=z, cond_w=cw)
                ms = timed(sstep, a.sample_steps, 3)
                out["sampler_step"].setdefault(v, {})[tag] = {"ms_per_step": ms, "img_s_1000_steps": n / (ms * 1000 / 1e3)}
        del m
        torch.cuda.empty_cache()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
