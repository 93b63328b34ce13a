"""-m gpu: `G.dropout > 0` — nn.Dropout(p) of ResBlock.out_layers (simple_unet.py:168-172) fused into the GroupNorm kernels.

The device draws its masks from a counter hash (b200ddpm.h: dropout_p / dropout_seed / dropout_step), not from torch's
Philox stream, so parity with the reference is "same function of the same mask": the tests rebuild the mask on the host
with the same integer arithmetic (`keep_multiplier`), hand it to torch / the oracle (`drop=`), and compare — kernel level
(forward, backward, every kernel family: fp32 shared-memory, bf16-source, streaming) and model level (loss and gradients
of a train step, eval mode untouched).
"""
import zlib

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import ddpm_oracle as O
from tests.gpu_util import ptr, rel, stream

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
M32 = 0xFFFFFFFF


def fmix32(h):
    h = h.astype(np.uint64) & M32
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & M32
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & M32
    h ^= h >> 16
    return h


def keep_multiplier(seed, step, N, HW, C, p):
    """[N, HW, C] float32 multiplier keep / (1 - p) exactly as gn.cu's drop_seed / drop_apply compute it."""
    thr = min(int(p * 65536.0 + 0.5), 65535)
    s = fmix32(np.uint64(seed) ^ fmix32(np.array((0x9E3779B9 * step + 0x7F4A7C15) & M32, dtype=np.uint64)))
    e4 = np.arange(N * HW * C // 4, dtype=np.uint64)
    a = fmix32((e4 * 0x9E3779B1 + s) & M32)
    b = fmix32((a + 0x632BE5AB) & M32)
    bits = np.stack([a & 0xFFFF, a >> 16, b & 0xFFFF, b >> 16], -1).reshape(N, HW, C)
    scale = np.float32(65536.0) / np.float32(65536 - thr)
    return torch.from_numpy(np.where(bits >= thr, scale, np.float32(0)).astype(np.float32)), thr


@pytest.fixture(scope="module")
def lib(built_lib):
    built_lib.load().b200_init(0)
    return built_lib


@pytest.mark.parametrize("N,HW,C2,src", [(3, 784, 0, "f32"), (2, 196, 128, "f32"), (2, 4096, 0, "f32"), (3, 784, 0, "bf16")])
def test_gn_dropout_forward_backward(lib, N, HW, C2, src):
    p, seed, step_v = 0.3, 0x1234ABCD, 5
    g = torch.Generator().manual_seed(HW + C2)
    Cc = 128 + C2
    x = torch.randn(N, HW, Cc, generator=g) * 2 + 0.5
    dy = torch.randn(N, HW, Cc, generator=g)
    if src == "bf16":
        x, dy = x.bfloat16().float(), dy.bfloat16().float()
    gamma, beta = 1 + 0.2 * torch.randn(Cc, generator=g), 0.1 * torch.randn(Cc, generator=g)
    mult, thr = keep_multiplier(seed, step_v, N, HW, Cc, p)
    assert abs((mult > 0).float().mean().item() - (1 - thr / 65536)) < 4 * (0.21 / (N * HW * Cc)) ** 0.5 + 1e-3  # keep rate
    # torch: nn.GroupNorm -> SiLU -> dropout with this mask
    xt = x.permute(0, 2, 1).contiguous().requires_grad_(True)
    gt, bt = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    yt = F.silu(F.group_norm(xt, 32, gt, bt, eps=1e-5)) * mult.permute(0, 2, 1)
    yt.backward(dy.permute(0, 2, 1))
    # device
    bfs = src == "bf16"
    xd = (x.bfloat16() if bfs else x).to(DEV)
    x1 = xd[..., :128].contiguous()
    x2 = xd[..., 128:].contiguous() if C2 else None
    dyd = (dy.bfloat16() if bfs else dy).to(DEV)
    gd, bd = gamma.to(DEV), beta.to(DEV)
    step = torch.tensor([step_v], dtype=torch.int32, device=DEV)
    stats = torch.zeros(N, 64, device=DEV)
    y = torch.full((N, HW, Cc), float("nan"), device=DEV)
    a = lib.GnArgs(x=ptr(x1), x2=ptr(x2), stats=ptr(stats), gamma=ptr(gd), beta=ptr(bd), y=ptr(y), N=N, HW=HW, C1=128, C2=C2,
                   x_ld=0, x2_ld=0, eps=1e-5, silu=1, round_tf32=0, y_dtype=0, x_dtype=1 if bfs else 0,
                   dropout_p=p, dropout_seed=seed, dropout_step=ptr(step))
    lib.call("b200_gn_fwd", a, stream())
    tol = 2e-3 if bfs else 1e-5  # bf16 sources use the one-MUFU sigmoid (2.5e-4 absolute)
    assert rel(y.cpu(), yt.detach().permute(0, 2, 1)) < tol
    # the same call with the counter advanced draws another mask; p = 0 leaves the fields unused
    step.fill_(step_v + 1)
    y2 = torch.empty_like(y)
    a.y = ptr(y2)
    lib.call("b200_gn_fwd", a, stream())
    assert ((y2 == 0) != (y == 0)).float().mean().item() > 0.2
    step.fill_(step_v)
    a.y = ptr(y)
    # backward
    dx = torch.full((N, HW, Cc), float("nan"), device=DEV)
    dgam, dbet = torch.zeros(Cc, device=DEV), torch.zeros(Cc, device=DEV)
    ws = torch.zeros(N * 64, device=DEV)
    b = lib.GnBwdArgs(fwd=a, dy=ptr(dyd), dx=ptr(dx), add1=None, add2=None, add1_ld=0, add2_ld=0, add_c=0, dgamma=ptr(dgam),
                      dbeta=ptr(dbet), dy_dtype=1 if bfs else 0)
    lib.call("b200_gn_bwd", b, stream(), ws=ptr(ws), ws_bytes=ws.numel() * 4)
    btol = 3e-3 if bfs else 1e-4
    assert rel(dx.cpu(), xt.grad.permute(0, 2, 1)) < btol
    assert rel(dgam.cpu(), gt.grad) < btol and rel(dbet.cpu(), bt.grad) < btol


def test_gn_dropout_rejects_bad_probability(lib):
    t = torch.zeros(1, 49, 128, device=DEV)
    st, gb = torch.zeros(1, 64, device=DEV), torch.ones(128, device=DEV)
    a = lib.GnArgs(x=ptr(t), x2=None, stats=ptr(st), gamma=ptr(gb), beta=ptr(gb), y=ptr(torch.empty_like(t)), N=1, HW=49,
                   C1=128, C2=0, x_ld=0, x2_ld=0, eps=1e-5, silu=1, round_tf32=0, dropout_p=1.0)
    with pytest.raises(lib.B200Error, match="dropout_p"):
        lib.call("b200_gn_fwd", a, stream())


@pytest.mark.parametrize("prec", ["tf32", "bf16"])
def test_train_step_with_dropout_vs_oracle(built_lib, prec):
    """The UNet with G.dropout = 0.2 — forward, all gradients, the training loss — against the oracle given the same masks;
    model.eval() switches the dropout off (sampling is bit-identical to a dropout-free model with the same weights)."""
    from generative_models_b200.model import DiffusionModelB200, default_G
    from generative_models_b200.unet_spec import RESBLOCKS
    p, B = 0.2, 4
    torch.manual_seed(0)
    m = DiffusionModelB200(default_G(device=DEV, mean_type="eps", sampler="noisy", timesteps=3, precision=prec, dropout=p,
                                     cf_drop_prob=0.0, graph=0))
    sd0 = O.random_state_dict(1)
    m.net.load_state_dict(sd0)
    g = torch.Generator().manual_seed(11)
    x = torch.rand(B, 1, 28, 28, generator=g) * 2 - 1
    y = torch.tensor([3, 1, 4, 1])
    eps = torch.randn(B, 1, 28, 28, generator=g)
    u = torch.rand(B, generator=g) * 0.8 + 0.1
    sizes = {"down.seq.1": 28, "down.seq.2": 28, "down.seq.4": 14, "down.seq.5": 14, "turn": 7, "up.seq.0.0": 7, "up.seq.1": 14,
             "up.seq.2": 14, "up.seq.3.0": 14, "up.seq.4": 28, "up.seq.5": 28, "up.seq.6": 28}
    assert set(sizes) == set(RESBLOCKS)
    step = int(m.adam_step.item())
    drop = {}
    for name, S in sizes.items():
        if name == "up.seq.3.0":  # built without the dropout argument in the reference (simple_unet.py:137-139)
            continue
        seed = (zlib.crc32((name + ".out_layers.0").encode()) ^ m.drop_seed) & M32
        mult, _ = keep_multiplier(seed, step, B, S * S, 128, p)
        drop[name] = mult.view(B, S, S, 128).permute(0, 3, 1, 2)
    # (1) the network itself: forward and every parameter gradient of a LINEAR loss sum(out * r) — the training loss
    # clips x~, and one clip decision flipped by TF32 noise moves the max-norm of a weight gradient by per cents, mask
    # or no mask — through the plan, as tests/test_gpu_attention.py does
    from generative_models_b200.plan import UNetPlan
    z = torch.randn(B, 1, 28, 28, generator=g)
    logsnr = torch.tensor([1.5, -3.0, 0.2, 4.0])
    r = torch.randn(B, 1, 28, 28, generator=g)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    out_ref = O.unet_forward(sd, z, logsnr, guide=y, drop=drop)
    (out_ref * r).sum().backward()
    out_nodrop = O.unet_forward(sd0, z, logsnr, guide=y)
    assert rel(out_nodrop, out_ref.detach()) > 1e-2  # the masks matter
    pl = UNetPlan(m.store, B, 28, train=True, bf16=(prec == "bf16"), dropout=p, drop_step=m.adam_step, drop_seed=m.drop_seed)
    pl.z_in.copy_(z.reshape(B, -1)); pl.t.copy_(logsnr); pl.guide.copy_(y)
    st = m._stream()
    m.store.grad.zero_()
    pl.forward(st)
    ntol, gtol = (1e-3, 6e-3) if prec == "tf32" else (8e-3, 4e-2)
    e = ((pl.out.cpu().view(B, 1, 28, 28).double() - out_ref.detach().double()).pow(2).mean().sqrt()
         / out_ref.detach().double().pow(2).mean().sqrt()).item()
    assert e < ntol, e
    pl.dout.copy_(r.reshape(B, -1))
    pl.backward(st)
    worst = ("", 0.0)
    for k, prm in m.net.named_parameters():
        if sd[k].grad is None:
            continue
        worst = max(worst, (k, rel(prm.grad.cpu(), sd[k].grad)), key=lambda t: t[1])
    assert worst[1] < gtol, worst
    del pl
    # (2) the training loss of the public API with the same masks
    ref_loss, _, _ = O.training_loss(sd0, x, y, eps, u, "eps", drop=drop)
    ref_nodrop, _, _ = O.training_loss(sd0, x, y, eps, u, "eps")
    assert abs(ref_loss.item() - ref_nodrop.item()) > 1e-3 * abs(ref_nodrop.item())
    m.store.grad.zero_()
    loss, _ = m.loss_and_grads(x.to(DEV), y.to(DEV), eps=eps.to(DEV), u=u.to(DEV))
    ltol = 3e-3 if prec == "tf32" else 3e-2
    assert abs(loss.item() - ref_loss.item()) < ltol * abs(ref_loss.item()), (loss.item(), ref_loss.item(), ref_nodrop.item())
    # the optimiser step advances the counter: the next step draws other masks
    m.train_step(x.to(DEV), y.to(DEV).clone(), eps=eps.to(DEV), u=u.to(DEV))
    assert int(m.adam_step.item()) == step + 1
    # eval mode: no dropout anywhere (a dropout-free model with the same weights samples the same bits)
    m.eval()
    torch.manual_seed(5)
    a = m.sample(2, y[:2].to(DEV))
    m0 = DiffusionModelB200(default_G(device=DEV, mean_type="eps", sampler="noisy", timesteps=3, precision=prec, cf_drop_prob=0.0))
    m0.load_state_dict(m.state_dict())
    m0.eval()
    torch.manual_seed(5)
    assert torch.equal(a, m0.sample(2, y[:2].to(DEV)))
    # and the evaluation loss ignores G.dropout
    l_eval, _ = m.loss(x.to(DEV), y.to(DEV))
    assert torch.isfinite(l_eval)
