"""Teacher-forced parity of the 1000-step sampler at the size bench.py times (n = 512), and the reference's own
TF32-vs-fp32 spread on the same device, inputs and weights ("state the norm — max|d| / max|ref|").

Shared by tests/test_gpu_sampler_parity.py (the bars) and tools/sampler_parity.py (the full table for DESIGN.md).

Checker: oracle/ddpm_oracle.py — pinned to the unmodified reference by tests/golden — run ON THE GPU in true fp32
(`cudnn.allow_tf32 = False`), i.e. the reference's own torch ops at full precision.  "ref_tf32" is the same oracle with
torch's default `cudnn.allow_tf32 = True`: what the reference itself computes on this box in fp32 storage.  Its distance
from the fp32 result is the reference's own spread; ours is held against the same fp32 result.

One step at loop index i (gaussian_diffusion.py:215-243 / :189-213 / :174-187, loop :287-292): z_t is a forward-diffused
synthetic image at logsnr_t (q_sample of uniform[-1,1] data with seeded noise — what the sampler sees at that index),
noise is seeded, labels cycle 0..9, guidance weights are 4*rand(n).
"""
import torch

from oracle import ddpm_oracle as O


def maxrel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()


def rmsrel(a, b):
    return ((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt().clamp_min(1e-30)).item()


def default_indices(T):
    """>= 20 loop indices: both ends and a spread."""
    base = {T - 1, T - 2, T - 3, 1, 0, 2}
    base |= {int(round(f * (T - 1))) for f in (0.97, 0.94, 0.9, 0.85, 0.8, 0.7, 0.6, 0.5, 0.4, 0.3, 0.2, 0.15, 0.1, 0.06, 0.03, 0.01)}
    return sorted((i for i in base if 0 <= i < T), reverse=True)


def oracle_step(sd, z, y, cond_w, lt, ls, noise, sampler, mean_type, allow_tf32):
    """x~, eps~, posterior mean (noisy) and z_s of one reference step on z's device."""
    old = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False  # torch default (the reference never changes it)
    try:
        with torch.no_grad():
            n = z.shape[0]
            ltv = lt.expand(n)
            out = O.unet_forward(sd, z, ltv, guide=y)
            xc, ec = O.run_model(out, z, ltv, mean_type)
            if cond_w is not None:
                out_u = O.unet_forward(sd, z, ltv, guide=-torch.ones_like(y))
                _, eu = O.run_model(out_u, z, ltv, mean_type)
                xc, ec = O.cf_guidance(ec, eu, z, ltv, cond_w)
            if sampler == "noisy":
                zs, mean, _ = O.posterior_noisy(z, xc, lt, ls, noise)
            else:
                zs = O.posterior_ddim(xc, ec, ls)
                mean = zs
            return {"out": out, "x": xc, "eps": ec, "mean": mean, "z_s": zs}
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def run(model, sd_dev, n, T, sampler, guided, indices=None, seed=0, mean_type="eps"):
    """Per-index rows {i, logsnr_t, ours_*: max-norm error vs fp32 oracle, ref_*: the reference's own TF32 spread}."""
    from generative_models_b200.schedule import sampler_table

    dev = model.device
    g = torch.Generator().manual_seed(seed)
    x0 = (torch.rand(n, 1, 28, 28, generator=g) * 2 - 1).to(dev)
    y = (torch.arange(n) % 10).to(dev)
    cw = (4.0 * torch.rand(n, generator=g)).to(dev) if guided else None
    table = sampler_table(T, dev)
    b = model._plan(n, "sample_g" if guided else "sample")
    rows = []
    for i in (indices or default_indices(T)):
        lt, ls = table[i, 0], table[i, 1]
        e0 = torch.randn(n, 1, 28, 28, generator=g).to(dev)
        noise = torch.randn(n, 1, 28, 28, generator=g).to(dev)
        z_t = O.q_sample(x0, lt.expand(n), e0)
        ref = oracle_step(sd_dev, z_t, y, cw, lt, ls, noise, sampler, mean_type, allow_tf32=False)
        rtf = oracle_step(sd_dev, z_t, y, cw, lt, ls, noise, sampler, mean_type, allow_tf32=True)
        x_o, e_o, zs_o = b.teacher_step(z_t, y, cw, i, noise if sampler == "noisy" else None)
        if sampler == "noisy":
            mean_o = b.teacher_step(z_t, y, cw, i, None)[2]  # zero noise: z_next is the posterior mean itself
        else:
            mean_o = zs_o
        want_z = ref["x"] if i == 0 else ref["z_s"]          # gaussian_diffusion.py:292
        want_m = ref["x"] if i == 0 else ref["mean"]
        rtf_z = rtf["x"] if i == 0 else rtf["z_s"]
        rtf_m = rtf["x"] if i == 0 else rtf["mean"]
        # raw net output of OUR conditional forward at this index (rows [0, n) of plan.out)
        out_o = b.plan.out[:n].view(n, 1, 28, 28)
        rows.append({
            "i": i, "logsnr_t": float(lt),
            "ours_out": maxrel(out_o, ref["out"]), "ref_out": maxrel(rtf["out"], ref["out"]),
            "ours_out_rms": rmsrel(out_o, ref["out"]), "ref_out_rms": rmsrel(rtf["out"], ref["out"]),
            "ours_x": maxrel(x_o, ref["x"]), "ref_x": maxrel(rtf["x"], ref["x"]),
            "ours_mean": maxrel(mean_o, want_m), "ref_mean": maxrel(rtf_m, want_m),
            "ours_z": maxrel(zs_o, want_z), "ref_z": maxrel(rtf_z, want_z),
        })
    return rows
