#!/usr/bin/env python
"""Developer tool (GPU): bf16-operand mode vs the TF32 mode and the reference fixture — errors and step times.

    python tools/bf16_check.py [--batch 512] [--steps 10]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from generative_models_b200.model import DiffusionModelB200, default_G  # noqa: E402
from generative_models_b200.plan import UNetPlan  # noqa: E402
from oracle import ddpm_oracle as O  # noqa: E402

DEV = "cuda:0"


def rms_rel(a, b):
    return ((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt().clamp_min(1e-30)).item()


def rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()


def model(precision, **over):
    m = DiffusionModelB200(default_G(device=DEV, precision=precision, **over))
    m.net.load_state_dict(O.random_state_dict(1))
    return m


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    golden = torch.load(os.path.join(ROOT, "tests", "golden", "reference_fixtures.pt"))
    outs = {}
    for prec in ("tf32", "bf16"):
        m = model(prec, mean_type="eps")
        p = UNetPlan(m.store, 3, 28, train=False, bf16=m.bf16)
        p.z_in.copy_(golden["unet_z"].reshape(3, -1))
        p.t.copy_(golden["unet_logsnr"])
        p.guide.copy_(golden["unet_guide"])
        p.forward(m._stream())
        o = p.out.cpu().view(3, 1, 28, 28)
        outs[prec] = o
        print(f"[{prec}] net out vs reference fixture: rms-rel {rms_rel(o, golden['unet_out']):.3e}  max-rel {rel(o, golden['unet_out']):.3e}")
    # gradients of a linear loss: bf16 vs tf32 vs oracle
    g = torch.Generator().manual_seed(77)
    B = 6
    x = torch.rand(B, 1, 28, 28, generator=g) * 2 - 1
    y = torch.tensor([0, 5, -1, 9, 3, -1])
    eps = torch.randn(B, 1, 28, 28, generator=g)
    u = torch.rand(B, generator=g) * 0.8 + 0.1
    r = torch.randn(B, 1, 28, 28, generator=g)
    sd0 = O.random_state_dict(1)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    _, _, aux = O.training_loss(sd, x, y, eps, u, "eps")
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    (O.unet_forward(sd, aux["z"], aux["logsnr"], guide=y) * r).sum().backward()
    for prec in ("tf32", "bf16"):
        m = model(prec, mean_type="eps")
        m.store.grad.zero_()
        _, b = m.loss_and_grads(x.to(DEV), y.to(DEV), eps=eps.to(DEV), u=u.to(DEV))
        print(f"[{prec}] net out (B=6) vs oracle: rms-rel {rms_rel(b.plan.out.cpu().view(B,1,28,28), aux['out'].detach()):.3e}")
        m.store.grad.zero_()
        st = m._stream()
        b.plan.forward(st)
        b.plan.dout.copy_(r.reshape(B, -1))
        b.plan.backward(st)
        errs = []
        for k, p in m.net.named_parameters():
            if sd[k].grad is None:
                continue
            errs.append((rel(p.grad.cpu(), sd[k].grad), rms_rel(p.grad.cpu(), sd[k].grad), k))
        errs.sort()
        print(f"[{prec}] linear-loss grads vs oracle autograd: worst max-rel {errs[-1][0]:.3e} ({errs[-1][2]}), "
              f"median {errs[len(errs)//2][0]:.3e}; worst rms-rel {max(e[1] for e in errs):.3e}")
    # step times
    B = a.batch
    xb = (torch.rand(B, 1, 28, 28, device=DEV) * 2 - 1)
    yb = torch.randint(0, 10, (B,), device=DEV)
    for prec in ("tf32", "bf16"):
        m = model(prec, mean_type="eps", sampler="noisy", timesteps=1000)
        for _ in range(4):
            out = m.train_step(xb, yb.clone())
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            out = m.train_step(xb, yb.clone())
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        print(f"[{prec}] train step B={B}: {ms:.2f} ms  {B / ms * 1e3:.0f} img/s  loss {out['loss'].item():.4f}  "
              f"act {m._plan(B, 'train').plan.activation_bytes() / 1e9:.2f} GB")
        # sampler: time 20 steps of the loop
        bnd = m._plan(B, "sample")
        bnd.begin_sampling(torch.randn(B, 1, 28, 28, device=DEV), yb, None)
        for _ in range(3):
            bnd.run_graphed(bnd.enqueue_sampler_step, "sampler_step")
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            bnd.run_graphed(bnd.enqueue_sampler_step, "sampler_step")
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"[{prec}] sampler step B={B}: {ms:.3f} ms  -> {B / ms:.1f} img/s at 1000 steps")
        del m, bnd
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
