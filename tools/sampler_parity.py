#!/usr/bin/env python
"""Print (and optionally save as JSON) the per-index parity table of the 1000-step sampler at n = 512: this repo vs the
fp32 oracle, beside the reference's own cuDNN-TF32-vs-fp32 spread on the same box (tests/sampler_parity_impl.py)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--T", type=int, default=1000)
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import __graft_entry__ as ge
    ge.build()
    from generative_models_b200.model import DiffusionModelB200, default_G
    from oracle import ddpm_oracle as O
    from tests import sampler_parity_impl as P
    res = {}
    for sampler, guided in (("noisy", False), ("noisy", True), ("ddim", False)):
        m = DiffusionModelB200(default_G(device="cuda:0", mean_type="eps", sampler=sampler, timesteps=a.T, graph=0,
                                         precision=a.precision))
        sd = O.random_state_dict(1)
        m.net.load_state_dict(sd)
        rows = P.run(m, {k: v.cuda() for k, v in sd.items()}, a.n, a.T, sampler, guided)
        res[f"{sampler}{'_guided' if guided else ''}"] = rows
        print(f"== {sampler} guided={guided}  n={a.n} T={a.T} precision={a.precision}   (ours / reference-TF32, max-norm vs fp32)")
        print("    i  logsnr_t   out(max)            out(rms)            x~                  post.mean           z_s")
        for r in rows:
            print(f" {r['i']:4d} {r['logsnr_t']:+8.2f}   {r['ours_out']:.2e}/{r['ref_out']:.2e}   {r['ours_out_rms']:.2e}/{r['ref_out_rms']:.2e}"
                  f"   {r['ours_x']:.2e}/{r['ref_x']:.2e}   {r['ours_mean']:.2e}/{r['ref_mean']:.2e}   {r['ours_z']:.2e}/{r['ref_z']:.2e}")
        del m
        torch.cuda.empty_cache()
    if a.out:
        json.dump(res, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
