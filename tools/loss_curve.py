#!/usr/bin/env python
"""Loss curve of the B200 path against the CPU oracle over many optimiser steps (north star: "loss curves tracking
within tolerance over 1k steps").

Both sides start from the same weights and see, step by step, the same batch, labels (some -1), eps and u; the oracle
side is `oracle/ddpm_oracle.py` (pinned to the reference) + autograd + torch.optim.Adam on the host cores, i.e. exactly
gms/diffusion/diffusion_model.py:63-74 without the autocast.  Single steps agree to ~1e-3 (TF32); over hundreds of Adam
steps the two weight trajectories separate chaotically, so the comparison is on the window-averaged curve.

    python tools/loss_curve.py --steps 1000 --batch 8 [--precision bf16] [--out profiles/r01_loss_curve.json]
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make_data(n_images, steps, B, seed=0):
    g = torch.Generator().manual_seed(seed)
    # a small "dataset" of smooth blobs so that there is something to learn (pure noise images have a flat loss)
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, 28), torch.linspace(-1, 1, 28), indexing="ij")
    labels = torch.randint(0, 10, (n_images,), generator=g)
    cx = (labels.float() / 9 * 1.2 - 0.6)[:, None, None] + 0.1 * torch.randn(n_images, 1, 1, generator=g)
    cy = 0.3 * torch.randn(n_images, 1, 1, generator=g)
    imgs = (torch.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / 0.08) * 2 - 1).unsqueeze(1)
    idx = torch.stack([torch.randperm(n_images, generator=g)[:B] for _ in range(steps)])
    drop = torch.rand(steps, B, generator=g) < 0.1
    eps = torch.randn(steps, B, 1, 28, 28, generator=g)
    u = torch.rand(steps, B, generator=g)
    return imgs, labels, idx, drop, eps, u


def smooth(v, w):
    t = torch.tensor(v, dtype=torch.float64)
    return torch.nn.functional.avg_pool1d(t[None, None], w, stride=w)[0, 0]


def run(steps, B, precision="tf32", lr=3e-4, window=50, dev="cuda:0", threads=None, verbose=True):
    from generative_models_b200.model import DiffusionModelB200, default_G
    from oracle import ddpm_oracle as O

    if threads:
        torch.set_num_threads(threads)
    imgs, labels, idx, drop, eps, u = make_data(256, steps, B)
    torch.manual_seed(0)
    m = DiffusionModelB200(default_G(device=dev, mean_type="eps", lr=lr, cf_drop_prob=0.0, precision=precision))
    sd = {k[4:]: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    opt = torch.optim.Adam(list(sd.values()), lr=lr)
    ours, ref = [], []
    t0 = time.time()
    for s in range(steps):
        x = imgs[idx[s]]
        y = labels[idx[s]].clone()
        y[drop[s]] = -1
        out = m.train_step(x.to(dev), y.to(dev), eps=eps[s].to(dev), u=u[s].to(dev))
        ours.append(out["loss"].item())
        opt.zero_grad()
        loss, _, _ = O.training_loss(sd, x, y, eps[s], u[s], "eps")
        loss.backward()
        opt.step()
        ref.append(loss.item())
        if verbose and (s + 1) % 100 == 0:
            print(f"step {s + 1}: ours {sum(ours[-100:]) / 100:.5f}  oracle {sum(ref[-100:]) / 100:.5f}  "
                  f"({time.time() - t0:.0f} s)", flush=True)
    so, sr = smooth(ours, window), smooth(ref, window)
    dev_rel = ((so - sr).abs() / sr).tolist()
    first = [abs(a - b) / abs(b) for a, b in zip(ours[:5], ref[:5])]
    return {"steps": steps, "batch": B, "lr": lr, "precision": precision, "window": window,
            "first5_rel_dev": first, "windowed_rel_dev_max": max(dev_rel), "windowed_rel_dev": dev_rel,
            "ours_windowed": so.tolist(), "oracle_windowed": sr.tolist(),
            "ours_first_last": [ours[0], sum(ours[-window:]) / window], "oracle_first_last": [ref[0], sum(ref[-window:]) / window]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--window", type=int, default=50)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import __graft_entry__ as ge
    ge.build()
    r = run(a.steps, a.batch, a.precision, window=a.window)
    print(json.dumps({k: v for k, v in r.items() if not k.endswith("windowed") and k != "windowed_rel_dev"}))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(r, f)


if __name__ == "__main__":
    main()
