"""Developer diagnostic: where the end-to-end train step (pinned host batch in, loss out, every step) spends its time
beyond the device-resident step."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from generative_models_b200.model import DiffusionModelB200, default_G  # noqa: E402


def main():
    dev = "cuda:0"
    B, K = 512, 20
    m = DiffusionModelB200(default_G(device=dev, mean_type="eps", sampler="noisy", timesteps=1000))
    host = [((torch.rand(B, 1, 28, 28) * 2 - 1).pin_memory(), torch.randint(0, 10, (B,)).pin_memory()) for _ in range(4)]
    devb = [(x.to(dev), y.to(dev)) for x, y in host]
    for i in range(40):
        m.train_step(devb[i % 4][0], devb[i % 4][1].clone())
    torch.cuda.synchronize()

    def run(mode):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        cpu = {"h2d": 0.0, "step": 0.0, "item": 0.0}
        e0.record()
        for i in range(K):
            t0 = time.perf_counter()
            if mode in ("e2e", "h2d_only"):
                x = host[i % 4][0].to(dev, non_blocking=True)
                y = host[i % 4][1].to(dev, non_blocking=True)
            else:
                x, y = devb[i % 4][0], devb[i % 4][1].clone()
            t1 = time.perf_counter()
            out = m.train_step(x, y)["loss"]
            t2 = time.perf_counter()
            if mode in ("e2e", "item_only"):
                out.item()
            t3 = time.perf_counter()
            cpu["h2d"] += t1 - t0
            cpu["step"] += t2 - t1
            cpu["item"] += t3 - t2
        e1.record()
        torch.cuda.synchronize()
        print(f"{mode:10s} {e0.elapsed_time(e1) / K:7.3f} ms/step   cpu per step: " +
              ", ".join(f"{k} {v / K * 1e3:.3f} ms" for k, v in cpu.items()), flush=True)

    for mode in ("resident", "e2e", "item_only", "h2d_only", "resident", "e2e"):
        run(mode)


if __name__ == "__main__":
    main()
