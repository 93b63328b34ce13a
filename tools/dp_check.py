#!/usr/bin/env python
"""Data-parallel parity on real GPUs (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/dp_check.py

Every rank runs `train_step` on its shard of one global batch (same weights, injected eps / u, no label drop) through the
product's data-parallel path — Up-bucket allreduce overlapped with the Down-path backward, then the rest, then the fused
Adam with grad_scale = 1/world — for 3 steps (eager warm-up, capture, replay).  Rank 0 then repeats the same 3 steps on
the WHOLE batch alone (world-size-1 path, one graph) and compares the summed gradient times 1/world against the
global-batch gradient.  The learning rate is 0 in both runs: the first Adam steps move every weight by lr * sign(g), so
with lr > 0 a 1e-6 difference in a near-zero gradient flips a weight by 2 lr and the later steps are no longer
comparable (the first version of this tool measured exactly that: 4e-3 weight differences after 3 steps at lr 1e-3);
the Adam moments still go through the reduced gradient and are compared too.  Prints one JSON line; exit code 1 on
mismatch.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from generative_models_b200 import dist as D
    rank, world, dev = D.init_from_env()
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    from generative_models_b200.model import DiffusionModelB200, default_G
    from oracle import ddpm_oracle as O

    Bg = 64
    g = torch.Generator().manual_seed(3)
    x = (torch.rand(Bg, 1, 28, 28, generator=g) * 2 - 1)
    y = torch.randint(0, 10, (Bg,), generator=g)
    eps = torch.randn(Bg, 1, 28, 28, generator=g)
    u = torch.rand(Bg, generator=g) * 0.9 + 0.05
    sd = O.random_state_dict(1)

    def run(lo, hi, use_dist):
        m = DiffusionModelB200(default_G(device=str(dev), mean_type="eps", cf_drop_prob=0.0, lr=0.0))
        m.net.load_state_dict(sd)
        losses = []
        for _ in range(3):
            out = m.train_step(x[lo:hi].to(dev), y[lo:hi].to(dev), eps=eps[lo:hi].to(dev), u=u[lo:hi].to(dev))
            losses.append(out["loss"].item())
        torch.cuda.synchronize()
        return m.store.exp_avg.clone(), m.store.grad.clone(), losses

    lo, hi = D.shard_range(Bg, rank, world)
    w_dp, g_dp, l_dp = run(lo, hi, True)
    res = None
    if rank == 0:
        # the same three steps on the whole batch, without the process group's collectives
        real_is_init = dist.is_initialized
        dist.is_initialized = lambda: False
        torch.distributed.is_initialized = dist.is_initialized
        try:
            w_1, g_1, l_1 = run(0, Bg, False)
            # the same shards, one after the other on this GPU alone (same per-launch batch as the data-parallel ranks,
            # hence the same kernel selection and tile geometry): their gradient sum is what the allreduce must produce
            parts = [run(*D.shard_range(Bg, r, world), False) for r in range(world)]
            g_sum = sum(p_[1] for p_ in parts)
            m_sum = sum(p_[0] for p_ in parts)
        finally:
            dist.is_initialized = real_is_init
            torch.distributed.is_initialized = real_is_init
        rel = lambda a, b: ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()
        # store.grad after the step = the SUM over ranks of local-mean gradients; x 1/world = the global-batch mean gradient
        e_g = rel(g_dp / world, g_1)
        e_w = rel(w_dp, w_1)  # Adam's first moment after 3 steps: (1 - 0.9^3) * mean gradient
        res = {"world": world, "global_batch": Bg,
               # the collective path itself: allreduced gradient vs the sum of the same shards computed on one GPU
               "allreduce_vs_shard_sum_rel_err": rel(g_dp, g_sum),
               "exp_avg_vs_shard_sum_rel_err": rel(w_dp * world, m_sum),
               # against ONE launch over the whole batch (other kernel selection for the small Linears and other tile
               # boundaries: fp32 reordering noise of ~1e-7 that TF32-rounded stores and the clip mask amplify)
               "grad_rel_err": e_g, "exp_avg_rel_err_after_3_steps": e_w,
               "loss_local_shard": l_dp, "loss_global_batch": l_1}
        print(json.dumps(res), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if res is not None and not (res["allreduce_vs_shard_sum_rel_err"] < 1e-4 and res["exp_avg_vs_shard_sum_rel_err"] < 1e-4
                                and res["grad_rel_err"] < 2e-2):
        sys.exit(1)


if __name__ == "__main__":
    main()
