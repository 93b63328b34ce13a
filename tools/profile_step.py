"""One eager (no CUDA graph) train step + one sampler step at the bench workload, for ncu launch lists / captures.

    python tools/profile_step.py [--batch 512] [--warm 2]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from generative_models_b200.model import DiffusionModelB200, default_G  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--warm", type=int, default=2)
    ap.add_argument("--no-sample", action="store_true")
    ap.add_argument("--precision", default="tf32", choices=["tf32", "bf16"])
    a = ap.parse_args()
    dev = "cuda:0"
    torch.manual_seed(0)
    m = DiffusionModelB200(default_G(device=dev, mean_type="eps", sampler="noisy", timesteps=1000, graph=0,
                                     precision=a.precision))
    gen = torch.Generator().manual_seed(1)
    for k, p in m.net.named_parameters():
        if ".out_layers.3." in k:
            p.data.copy_((torch.randn(p.shape, generator=gen) * 0.02).to(dev))
    x = (torch.rand(a.batch, 1, 28, 28, generator=gen) * 2 - 1).to(dev)
    y = torch.randint(0, 10, (a.batch,), generator=gen).to(dev)
    for _ in range(a.warm + 1):
        loss = m.train_step(x, y.clone())["loss"]
    torch.cuda.synchronize()
    print("train loss", loss.item())
    if not a.no_sample:
        b = m._plan(a.batch, "sample")
        b.begin_sampling(torch.randn(a.batch, 1, 28, 28, device=dev), y, None)
        for _ in range(2):
            b.enqueue_sampler_step()
        torch.cuda.synchronize()
        print("sampler steps ok", torch.isfinite(b.plan.z_in).all().item())


if __name__ == "__main__":
    main()
