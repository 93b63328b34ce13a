"""Per-launch device times of one eager train step (CUDA events around every C-ABI call), with the geometry of each
tensor-core op: where a step's time goes, layer by layer.

    python tools/per_op.py [--batch 512] [--csv gpurun_out/per_op.csv]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from generative_models_b200.model import DiffusionModelB200, default_G  # noqa: E402


def describe(o):
    a = o.args
    if o.name == "b200_conv_igemm":
        return f"mode{a.mode} N{a.N} H{a.H} Ho{a.Ho} C{a.C1 + a.C2} R{a.R} s{a.stride} n_off{a.n_off}"
    if o.name == "b200_conv_wgrad":
        return f"N{a.N} H{a.H} Ho{a.Ho} C{a.C1 + a.C2} R{a.R} s{a.stride}"
    if o.name == "b200_gn_fwd":
        return f"N{a.N} HW{a.HW} C{a.C1 + a.C2}"
    if o.name == "b200_gn_bwd":
        return f"N{a.fwd.N} HW{a.fwd.HW} C{a.fwd.C1 + a.fwd.C2} add{int(bool(a.add1)) + int(bool(a.add2))}"
    if o.name == "b200_colsum":
        return f"N{a.N} HW{a.HW}"
    return ""


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--csv", default="")
    a = ap.parse_args()
    dev = "cuda:0"
    torch.manual_seed(0)
    m = DiffusionModelB200(default_G(device=dev, mean_type="eps", sampler="noisy", timesteps=1000, graph=0))
    gen = torch.Generator().manual_seed(1)
    for k, p in m.net.named_parameters():
        if ".out_layers.3." in k:
            p.data.copy_((torch.randn(p.shape, generator=gen) * 0.02).to(dev))
    x, y = bench.synthetic_batch(a.batch, 0)
    x, y = x.to(dev), y.to(dev)
    for _ in range(3):
        m.train_step(x, y.clone())
    b = m._plan(a.batch, "train")
    p = b.plan
    ops = [o for g in (b.zero_ops, b.pre_ops, p.refresh_ops, p.embed_ops[0], p.body_ops, p.head_ops[0], b.loss_ops,
                       p.bwd_ops, b.adam_ops) for o in g]
    st = torch.cuda.current_stream()
    sp = m._stream()
    best = [1e9] * len(ops)
    for _ in range(3):
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(len(ops) + 1)]
        evs[0].record(st)
        for i, o in enumerate(ops):
            assert o.fn(o.ref, o.ws, o.wsb, sp) == 0, o.name
            evs[i + 1].record(st)
        torch.cuda.synchronize()
        best = [min(bv, evs[i].elapsed_time(evs[i + 1])) for i, bv in enumerate(best)]
    rows = []
    for o, ms in zip(ops, best):
        fl = bench.conv_flops(o)
        rows.append((o.name, describe(o), ms * 1e3, fl / (ms / 1e3) / 1e12 if fl else 0.0))
    total = sum(r[2] for r in rows)
    lines = ["op,geometry,us,tflops"] + [f"{n},{d},{us:.1f},{tf:.1f}" for n, d, us, tf in rows]
    lines.append(f"total,,{total:.1f},")
    out = "\n".join(lines)
    print(out)
    if a.csv:
        open(a.csv, "w").write(out + "\n")


if __name__ == "__main__":
    main()
