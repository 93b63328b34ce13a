#!/usr/bin/env python
"""The three per-pixel diffusion kernels at a size where bandwidth matters (BASELINE.json configs[3]: n = 8192 images per
GPU): CUDA-event time and achieved HBM GB/s against their algorithmic bytes (q_sample 12 B/px, loss fwd+bwd 20 B/px,
ancestral step 16 B/px — SURVEY.md §8d), straight through the C ABI.  Also the target of the ncu `--set full` captures
kept under profiles/ (north star: "achieved HBM GB/s ... for the elementwise kernels, evidenced by ncu").

    python tools/profile_elementwise.py [--n 8192] [--reps 20]
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=8192)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import __graft_entry__ as ge
    ge.build()
    from generative_models_b200 import lib
    from generative_models_b200.schedule import sampler_table
    dev = "cuda:0"
    n, D = a.n, 784
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    x, eps, z, out, noise, dout = (torch.randn(n, D, device=dev) for _ in range(6))
    logsnr = torch.linspace(-10, 10, n, device=dev)
    loss_b, loss_mean = torch.zeros(n, device=dev), torch.zeros(1, device=dev)
    sched = sampler_table(1000, dev)
    step_dev = torch.tensor([500], dtype=torch.int32, device=dev)
    # a buffer larger than L2 written between repetitions would dominate the timing at these sizes; instead every tensor
    # here (25.7 MB each, 77-128 MB per kernel) is touched once per launch and the set of kernels cycles, so a launch
    # never finds its inputs in the 126 MB L2 left by its previous launch
    calls = {
        "q_sample_kernel": ("b200_q_sample", lib.QSampleArgs(x=p(x), eps=p(eps), logsnr=p(logsnr), z=p(z), B=n, D=D), 12),
        "loss_kernel": ("b200_loss", lib.LossArgs(out=p(out), z=p(z), x=p(x), eps=p(eps), logsnr=p(logsnr), loss_b=p(loss_b),
                                                  mse2=None, dout=p(dout), loss_mean=p(loss_mean), B=n, D=D, mean_type=0,
                                                  grad_scale=1.0 / n), 20),
        "sampler_step_kernel": ("b200_sampler_step", lib.SamplerStepArgs(z=p(z), out_c=p(out), out_u=None, cond_w=None,
                                                                         noise=p(noise), sched=p(sched), step_dev=p(step_dev),
                                                                         z_next=p(z), n=n, D=D, step=0, mean_type=0, sampler=1), 16),
    }
    flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
    res = {}
    for name, (fn, args, bpp) in calls.items():
        lib.call(fn, args, st)
        ts = []
        for _ in range(a.reps):
            flush.sum()  # evict with a READ pass (256 MB > 126 MB L2): a write pass leaves the L2 full of dirty lines whose
            # write-back then competes with the timed kernel's own traffic
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            lib.call(fn, args, st)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        med = ts[len(ts) // 2]
        res[name] = {"us": med * 1e3, "algorithmic_bytes": n * D * bpp, "GBps": n * D * bpp / med / 1e6}
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs")
    except Exception:
        pass
    for k, v in res.items():
        v["frac_of_measured_copy_peak"] = v["GBps"] / peak if peak else None
    print(json.dumps({"n": n, "D": D, "l2_flush": "256 MB read pass between repetitions", "hbm_peak_gbs": peak, "kernels": res}))


if __name__ == "__main__":
    main()
