"""CPU restatement of the DDPM hot path of matwilso/generative_models (gms/diffusion) — TEST INFRASTRUCTURE ONLY.

This module is the parity oracle: a plain-PyTorch (CPU, fp32 or fp64) functional restatement of the reference's
algorithm, written against a flat ``state_dict`` instead of the reference's nn.Module tree.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import it; the
product package (generative_models_b200/) never does and fails loudly without its CUDA library.

Pinning: the reference's own tests hold no golden vectors (tests/test_models.py:7-13 is an exit-code smoke
test), so this oracle is pinned against the reference modules executed live (oracle/make_golden.py imports
/root/reference, tests/test_oracle_vs_reference.py compares when that path exists) and against the fixtures
that script committed under tests/golden/.

Every function cites the reference file:line (relative to the reference root) it follows.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------- schedule
# gms/diffusion/diffusion_utils.py:198-201 (_logsnr_schedule_cosine) with gaussian_diffusion.py:33-35
# (logsnr_min=-20, logsnr_max=20).  a, b are numpy float64 scalars multiplied into an fp32 tensor, which
# torch evaluates in fp32 (python-scalar promotion), exactly like the reference.
LOGSNR_MIN, LOGSNR_MAX = -20.0, 20.0
_B = np.arctan(np.exp(-0.5 * LOGSNR_MAX))
_A = np.arctan(np.exp(-0.5 * LOGSNR_MIN)) - _B


def logsnr_cosine(t):
    return -2.0 * torch.log(torch.tan(_A * t + _B))


def bcast(v, like):
    """diffusion_utils.py:126-130 broadcast_from_left."""
    return v.reshape(v.shape + (1,) * (like.ndim - v.ndim)).expand(like.shape)


# ----------------------------------------------------------------------------- x / eps / v conversions
def x_from_eps(z, eps, logsnr):  # diffusion_utils.py:76-82
    l = bcast(logsnr, z)
    return torch.sqrt(1.0 + torch.exp(-l)) * (z - eps * torch.rsqrt(1.0 + torch.exp(l)))


def eps_from_x(z, x, logsnr):  # diffusion_utils.py:85-91
    l = bcast(logsnr, z)
    return torch.sqrt(1.0 + torch.exp(l)) * (z - x * torch.rsqrt(1.0 + torch.exp(-l)))


def x_from_v(z, v, logsnr):  # diffusion_utils.py:101-105
    l = bcast(logsnr, z)
    return torch.sqrt(torch.sigmoid(l)) * z - torch.sqrt(torch.sigmoid(-l)) * v


def q_sample(x, logsnr, eps):
    """diffusion_utils.py:65-73 (diffusion_forward) + gaussian_diffusion.py:99-100."""
    l = bcast(logsnr, x)
    return x * torch.sqrt(torch.sigmoid(l)) + torch.sqrt(torch.sigmoid(-l)) * eps


def run_model(out, z, logsnr, mean_type):
    """gaussian_diffusion.py:45-79 (_run_model) given the raw net output; returns (x_clipped, eps_from_clipped_x)."""
    if mean_type == "eps":
        xh = x_from_eps(z, out, logsnr)
    elif mean_type == "x":
        xh = out
    elif mean_type == "v":
        xh = x_from_v(z, out, logsnr)
    else:
        raise NotImplementedError(mean_type)
    xc = torch.clip(xh, -1.0, 1.0)
    return xc, eps_from_x(z, xc, logsnr)


# ----------------------------------------------------------------------------- network (simple_unet.py)
def sincos_embedding(t, dim, max_period):
    """simple_unet.py:205-224 timestep_embedding (cos first, then sin)."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32) / half).to(device=t.device, dtype=t.dtype)
    args = t[:, None] * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def _lin(sd, name, x):
    return F.linear(x, sd[name + ".weight"], sd[name + ".bias"])


def _conv(sd, name, x, stride=1, padding=1):
    return F.conv2d(x, sd[name + ".weight"], sd[name + ".bias"], stride=stride, padding=padding)


def _gn_silu(sd, name, x):
    return F.silu(F.group_norm(x, 32, sd[name + ".weight"], sd[name + ".bias"], eps=1e-5))


def resblock(sd, pre, x, emb, drop=None):
    """simple_unet.py:155-186.  `drop` (optional dict: block name -> [B, C, H, W] multiplier, i.e. keep / (1 - p)) stands
    for nn.Dropout(p) between the SiLU and the convolution of out_layers (:168-172) with an explicit mask, so that a test
    can hand the oracle the mask the device drew; None = p = 0 / eval mode."""
    h = _conv(sd, pre + ".in_layers.2", _gn_silu(sd, pre + ".in_layers.0", x))
    h = h + _lin(sd, pre + ".emb_layers.1", F.silu(emb))[:, :, None, None]
    a = _gn_silu(sd, pre + ".out_layers.0", h)
    if drop is not None and pre in drop:
        a = a * drop[pre]
    h = _conv(sd, pre + ".out_layers.3", a)
    if (pre + ".skip_connection.weight") in sd:
        x = _conv(sd, pre + ".skip_connection", x, padding=0)
    return x + h


def attention_block(sd, pre, x):
    """EXTENSION (no reference counterpart: simple_unet.py:16-72 has no attention; the north star / BASELINE.json
    configs[4] ask for one): guided-diffusion's single-head AttentionBlock — x + proj(softmax(q k^T / sqrt(C)) v) with
    q, k, v, proj 1x1 convolutions of GroupNorm(32, C)(x), proj zero-initialised.  "Parity unpinned": this function is the
    only checker."""
    B, C, H, W = x.shape
    h = F.group_norm(x, 32, sd[pre + ".norm.weight"], sd[pre + ".norm.bias"], eps=1e-5)
    q, k, v = (_conv(sd, f"{pre}.{n}", h, padding=0).reshape(B, C, H * W) for n in ("q", "k", "v"))
    w = torch.softmax(torch.einsum("bct,bcs->bts", q, k) / math.sqrt(C), dim=-1)
    a = torch.einsum("bts,bcs->bct", w, v).reshape(B, C, H, W)
    return x + _conv(sd, pre + ".proj_out", a, padding=0)


def embed(sd, logsnr, guide=None, cond_w=None):
    """simple_unet.py:45-64: time MLP (+ masked class-guide MLP) (+ cond_w MLP)."""
    e = _lin(sd, "time_embed.2", F.silu(_lin(sd, "time_embed.0", sincos_embedding(logsnr.float(), 64, 256))))
    if guide is not None:
        mask = guide == -1
        onehot = F.one_hot(torch.where(mask, torch.zeros_like(guide), guide), 10).to(e.dtype)
        ge = _lin(sd, "guide_embed.2", F.silu(_lin(sd, "guide_embed.0", onehot)))
        e = e + torch.where(mask[:, None], torch.zeros_like(ge), ge)
    if cond_w is not None:
        e = e + _lin(sd, "cond_w_embed.2", F.silu(_lin(sd, "cond_w_embed.0", sincos_embedding(cond_w, 64, 4))))
    return e


def unet_forward(sd, x, logsnr, guide=None, cond_w=None, taps=None, drop=None):
    """simple_unet.py:44-72 + Down :87-109, Up :125-152, out head :38-42.  `sd` keys have no 'net.' prefix.
    `taps` (optional dict) receives named intermediates for layer-level parity tests."""
    emb = embed(sd, logsnr, guide, cond_w)
    skips = []
    h = _conv(sd, "down.seq.0.conv", x)
    skips.append(h)
    for i in (1, 2):
        h = resblock(sd, f"down.seq.{i}", h, emb, drop)
        skips.append(h)
    h = _conv(sd, "down.seq.3.conv", h, stride=2)
    skips.append(h)
    for i in (4, 5):
        h = resblock(sd, f"down.seq.{i}", h, emb, drop)
        skips.append(h)
    h = _conv(sd, "down.seq.6.conv", h, stride=2)
    skips.append(h)
    if taps is not None:
        taps["emb"] = emb
        taps["down"] = h
    h = resblock(sd, "turn", h, emb, drop)
    if "attn.q.weight" in sd:  # extension: self-attention at the lowest-resolution level, after the turn block
        h = attention_block(sd, "attn", h)
    if taps is not None:
        taps["turn"] = h
    for i in range(7):
        h = torch.cat([h, skips.pop()], 1)
        if i in (0, 3):
            h = resblock(sd, f"up.seq.{i}.0", h, emb, drop)
            h = F.interpolate(h, scale_factor=2, mode="nearest")
            h = _conv(sd, f"up.seq.{i}.1.conv", h)
        else:
            h = resblock(sd, f"up.seq.{i}", h, emb, drop)
    if taps is not None:
        taps["up"] = h
    return _conv(sd, "out.2", _gn_silu(sd, "out.0", h))


# ----------------------------------------------------------------------------- training loss
def training_loss(sd, x, y, eps, u, mean_type="eps", drop=None):
    """gaussian_diffusion.py:81-100,156-172 (non-distillation) + diffusion_model.py:76-80.
    eps, u are the injected random draws (reference draws eps = randn first, then u = rand).
    Returns (mean loss, per-sample loss, dict of intermediates)."""
    logsnr = logsnr_cosine(u)
    z = q_sample(x, logsnr, eps)
    out = unet_forward(sd, z, logsnr, guide=y, drop=drop)
    xc, ec = run_model(out, z, logsnr, mean_type)
    x_mse = ((xc - x) ** 2).flatten(1).mean(1)
    e_mse = ((ec - eps) ** 2).flatten(1).mean(1)
    per = torch.maximum(x_mse, e_mse)
    return per.mean(), per, {"logsnr": logsnr, "z": z, "out": out, "x_mse": x_mse, "eps_mse": e_mse}


# ----------------------------------------------------------------------------- progressive distillation
def ddim_step(net, z_t, logsnr_t, logsnr_s, mean_type, cond_w=None):
    """gaussian_diffusion.py:189-213.  net(z, logsnr_vec, uncond) -> raw output; logsnr_t / logsnr_s are [B] vectors.
    Returns (z_s, x_pred, eps_pred)."""
    xc, ec = run_model(net(z_t, logsnr_t, False), z_t, logsnr_t, mean_type)
    if cond_w is not None:
        _, eu = run_model(net(z_t, logsnr_t, True), z_t, logsnr_t, mean_type)
        xc, ec = cf_guidance(ec, eu, z_t, logsnr_t, cond_w)
    return posterior_ddim(xc, ec, logsnr_s), xc, ec


def distill_targets(sd_teacher, z_t, y, logsnr, u, cond_w, mode, num_steps, mean_type, i=None):
    """gaussian_diffusion.py:105-154: the teacher's denoising targets (x_target, eps_target), no gradient.
    step1: one classifier-free-guided DDIM step of the (unconditioned-on-w) teacher; step2: two half DDIM steps of the
    w-conditioned teacher and the x implied by where they land."""
    with torch.no_grad():
        u_s = u - 1.0 / num_steps
        logsnr_s = logsnr_cosine(u_s)
        if mode == "step1":
            net = lambda z, l, uncond: unet_forward(sd_teacher, z, l, guide=(-torch.ones_like(y) if uncond else y))
            _, x_t, e_t = ddim_step(net, z_t, logsnr, logsnr_s, mean_type, cond_w=cond_w)
            return x_t, e_t
        assert mode == "step2"
        net = lambda z, l, uncond: unet_forward(sd_teacher, z, l, guide=y, cond_w=cond_w)
        logsnr_mid = logsnr_cosine(u - 0.5 / num_steps)
        z_mid, _, _ = ddim_step(net, z_t, logsnr, logsnr_mid, mean_type)
        z_teacher, x_pred_teacher, _ = ddim_step(net, z_mid, logsnr_mid, logsnr_s, mean_type)
        bc = lambda v: bcast(v, z_t)
        alpha_s = bc(torch.sqrt(torch.sigmoid(logsnr_s)))
        alpha_t = bc(torch.sqrt(torch.sigmoid(logsnr)))
        stdv_frac = bc(torch.exp(0.5 * (F.softplus(logsnr) - F.softplus(logsnr_s))))
        x_target = (z_teacher - stdv_frac * z_t) / (alpha_s - stdv_frac * alpha_t)
        x_target = torch.where(bc(i == 0), x_pred_teacher, x_target)
        return x_target, eps_from_x(z_t, x_target, logsnr)


def training_loss_distill(sd, sd_teacher, x, y, eps, u, cond_w, mode, num_steps, mean_type="eps", i=None):
    """gaussian_diffusion.py:81-172 with a teacher: the student is conditioned on cond_w (simple_unet.py:60-64), the
    targets come from the teacher, the loss is eps_mse ('snr', step1) or max(x_mse, eps_mse) ('snr_trunc', step2).
    step2 draws a discrete u = (i+1)/num_steps (:87-91): pass `i` (int64 [B]) and u is derived from it."""
    if mode == "step2":
        u = (i + 1).to(x.dtype) / num_steps
    logsnr = logsnr_cosine(u)
    z = q_sample(x, logsnr, eps)
    x_t, e_t = distill_targets(sd_teacher, z, y, logsnr, u, cond_w, mode, num_steps, mean_type, i=i)
    out = unet_forward(sd, z, logsnr, guide=y, cond_w=cond_w)
    xc, ec = run_model(out, z, logsnr, mean_type)
    x_mse = ((xc - x_t) ** 2).flatten(1).mean(1)
    e_mse = ((ec - e_t) ** 2).flatten(1).mean(1)
    per = e_mse if mode == "step1" else torch.maximum(x_mse, e_mse)
    return per.mean(), per, {"logsnr": logsnr, "z": z, "out": out, "x_target": x_t, "eps_target": e_t}


def loss_from_out(out, z, x, eps, logsnr, mean_type="eps"):
    """Same loss as training_loss but from a supplied raw net output (for the loss-kernel parity test)."""
    xc, ec = run_model(out, z, logsnr, mean_type)
    x_mse = ((xc - x) ** 2).flatten(1).mean(1)
    e_mse = ((ec - eps) ** 2).flatten(1).mean(1)
    return torch.maximum(x_mse, e_mse)


# ----------------------------------------------------------------------------- sampler
def cf_guidance(e_cond, e_uncond, z, logsnr, w):
    """gaussian_diffusion.py:174-187."""
    wb = bcast(w, z)
    e = (1 + wb) * e_cond + (-wb) * e_uncond
    xc = torch.clip(x_from_eps(z, e, logsnr), -1.0, 1.0)
    return xc, eps_from_x(z, xc, logsnr)


def posterior_noisy(z_t, x_pred, logsnr_t, logsnr_s, noise):
    """gaussian_diffusion.py:233-242 with diffusion_utils.py:34-51,62 (x_logvar='large')."""
    lt, ls = bcast(logsnr_t, z_t), bcast(logsnr_s, z_t)
    alpha_st = torch.sqrt((1.0 + torch.exp(-lt)) / (1.0 + torch.exp(-ls)))
    alpha_s = torch.sqrt(torch.sigmoid(ls))
    r = torch.exp(lt - ls)
    one_minus_r = -torch.expm1(lt - ls)
    mean = r * alpha_st * z_t + one_minus_r * alpha_s * x_pred
    std = torch.sqrt(one_minus_r * torch.sigmoid(-lt))
    return mean + std * noise, mean, std


def posterior_ddim(x_pred, eps_pred, logsnr_s):
    """gaussian_diffusion.py:210-212."""
    ls = bcast(logsnr_s, x_pred)
    return torch.sqrt(torch.sigmoid(ls)) * x_pred + torch.sqrt(torch.sigmoid(-ls)) * eps_pred


def sample_loop(net, init_z, num_steps, sampler="noisy", mean_type="eps", cond_w=None, noises=None):
    """gaussian_diffusion.py:245-296.  `net(z, logsnr_vec, uncond)` returns the raw net output; `noises[k]` is the
    k-th per-step normal draw (k counts up from the first executed step, i = T-1).  Returns the three histories."""
    z = init_z
    zs, xs, es = [], [], []
    n = init_z.shape[0]
    for k, i in enumerate(range(num_steps - 1, -1, -1)):
        ti = torch.tensor(i, device=init_z.device)  # gaussian_diffusion.py:288
        lt = logsnr_cosine((ti + 1.0) / num_steps)
        ls = logsnr_cosine(ti / num_steps)
        ltv = lt.expand(n)
        xc, ec = run_model(net(z, ltv, False), z, ltv, mean_type)
        if cond_w is not None:
            _, eu = run_model(net(z, ltv, True), z, ltv, mean_type)
            xc, ec = cf_guidance(ec, eu, z, ltv, cond_w)
        if sampler == "noisy":
            z_s, _, _ = posterior_noisy(z, xc, lt, ls, noises[k])
        elif sampler == "ddim":
            z_s = posterior_ddim(xc, ec, ls)
        else:
            raise NotImplementedError(sampler)
        z = xc if i == 0 else z_s
        zs.append(z)
        xs.append(xc)
        es.append(ec)
    return torch.stack(zs), torch.stack(xs), torch.stack(es)


# ----------------------------------------------------------------------------- helpers for tests / fixtures
def param_shapes(C=128, img_channels=1, attention=False):
    """Appendix C of SURVEY.md: the 160 tensors of SimpleUnet(C) (keys without the 'net.' prefix), in
    registration order (simple_unet.py:16-42,87-186)."""
    E = 2 * C
    out = []

    def lin(name, i, o):
        out.append((name + ".weight", (o, i)))
        out.append((name + ".bias", (o,)))

    def conv(name, i, o, k):
        out.append((name + ".weight", (o, i, k, k)))
        out.append((name + ".bias", (o,)))

    def gn(name, c):
        out.append((name + ".weight", (c,)))
        out.append((name + ".bias", (c,)))

    def rb(pre, cin, cout):
        gn(pre + ".in_layers.0", cin)
        conv(pre + ".in_layers.2", cin, cout, 3)
        lin(pre + ".emb_layers.1", E, cout)
        gn(pre + ".out_layers.0", cout)
        conv(pre + ".out_layers.3", cout, cout, 3)
        if cin != cout:
            conv(pre + ".skip_connection", cin, cout, 1)

    for nm, i in (("time_embed", 64), ("cond_w_embed", 64), ("guide_embed", 10)):
        lin(nm + ".0", i, E)
        lin(nm + ".2", E, E)
    conv("down.seq.0.conv", img_channels, C, 3)
    rb("down.seq.1", C, C)
    rb("down.seq.2", C, C)
    conv("down.seq.3.conv", C, C, 3)
    rb("down.seq.4", C, C)
    rb("down.seq.5", C, C)
    conv("down.seq.6.conv", C, C, 3)
    rb("turn", C, C)
    if attention:
        gn("attn.norm", C)
        for n in ("q", "k", "v", "proj_out"):
            conv("attn." + n, C, C, 1)
    for i in range(7):
        if i in (0, 3):
            rb(f"up.seq.{i}.0", 2 * C, C)
            conv(f"up.seq.{i}.1.conv", C, C, 3)
        else:
            rb(f"up.seq.{i}", 2 * C, C)
    gn("out.0", C)
    conv("out.2", C, img_channels, 3)  # img_channels = 3: the configs[4] extension (no reference counterpart)
    return out


def random_state_dict(seed, C=128, dtype=torch.float32, img_channels=1, attention=False):
    """Deterministic non-degenerate weights (the reference zero-inits every out_layers.3, simple_unet.py:172,
    which would hide conv2/GN2 bugs): fan-in scaled normals from a CPU generator, GN weights near 1."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape in param_shapes(C, img_channels, attention):
        if ".weight" in name and len(shape) == 1:  # GroupNorm gamma
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".bias"):
            t = 0.05 * torch.randn(shape, generator=g)
        else:
            fan_in = int(np.prod(shape[1:]))
            t = torch.randn(shape, generator=g) / math.sqrt(fan_in)
        sd[name] = t.to(dtype)
    return sd
