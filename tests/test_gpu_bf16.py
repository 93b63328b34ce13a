"""-m gpu parity of the bf16-operand kernels (BASELINE.json configs[2]): strip conv fprop / dgrad, wgrad, the bf16
outputs of GroupNorm forward / backward and the cast kernel, through the C ABI.

bf16 x bf16 products are exact in fp32 and the tensor cores accumulate in fp32, so against an fp32 reference computed
from the SAME bf16-rounded operands the only difference is summation order: tolerance 1e-4 of max|ref|.  (The bf16
rounding of the operands themselves — 2^-9 relative — is the model-level tolerance, tests/test_gpu_model.py.)"""
import pytest
import torch
import torch.nn.functional as F

from tests.gpu_util import nhwc, ptr, rel, stream, w_khwc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-4


@pytest.fixture(scope="module")
def lib(built_lib):
    built_lib.load().b200_init(0)
    return built_lib


def rn(g, *s):
    return torch.randn(*s, generator=g)


def bf(t):
    """round to bf16 and back (what the kernels see)"""
    return t.to(torch.bfloat16).float()


def dev_bf(t):
    return t.to(torch.bfloat16).contiguous().to(DEV)


def fprop_bf16(lib, x, w, pad, x2=None, bias=None, emb=None, residual=None):
    N, C1, H, W = x.shape
    C2 = 0 if x2 is None else x2.shape[1]
    R = w.shape[2]
    xh = dev_bf(nhwc(x))
    x2h = dev_bf(nhwc(x2)) if x2 is not None else None
    wk = dev_bf(w_khwc(w))
    out = torch.full((N, H, W, 128), float("nan"), device=DEV)
    resh = nhwc(residual).to(DEV) if residual is not None else None
    bd = bias.to(DEV) if bias is not None else None
    ed = emb.to(DEV) if emb is not None else None
    a = lib.ConvArgs(x=ptr(xh), x2=ptr(x2h), w=ptr(wk), out=ptr(out), bias=ptr(bd), emb=ptr(ed), residual=ptr(resh), N=N,
                     H=H, W=W, C1=C1, C2=C2, Ho=H, Wo=W, R=R, S=R, stride=1, pad=pad, mode=0, w_cout=128,
                     w_cin=C1 + C2, n_off=0, out_ld=128, emb_ld=128, res_ld=128, op_dtype=1)
    lib.call("b200_conv_igemm", a, stream())
    return out.permute(0, 3, 1, 2).cpu()


def dgrad_bf16(lib, dy, w, pad):
    N, Co, H, W = dy.shape
    Cin, R = w.shape[1], w.shape[2]
    dyh = dev_bf(nhwc(dy))
    wk = dev_bf(w_khwc(w))
    out = torch.full((N, H, W, Cin), float("nan"), device=DEV)
    for n_off in range(0, Cin, 128):
        a = lib.ConvArgs(x=ptr(dyh), x2=None, w=ptr(wk), out=ptr(out, n_off), bias=None, emb=None, residual=None, N=N, H=H,
                         W=W, C1=Co, C2=0, Ho=H, Wo=W, R=R, S=R, stride=1, pad=pad, mode=1, w_cout=Co, w_cin=Cin,
                         n_off=n_off, out_ld=Cin, emb_ld=0, res_ld=Cin, op_dtype=1)
        lib.call("b200_conv_igemm", a, stream())
    return out.permute(0, 3, 1, 2).cpu()


def wgrad_bf16(lib, x, dy, R, pad, x2=None):
    N, C1, H, W = x.shape
    C2 = 0 if x2 is None else x2.shape[1]
    xh = dev_bf(nhwc(x))
    x2h = dev_bf(nhwc(x2)) if x2 is not None else None
    dyh = dev_bf(nhwc(dy))
    dw = torch.zeros((128, R, R, C1 + C2), device=DEV)
    a = lib.WgradArgs(x=ptr(xh), x2=ptr(x2h), dy=ptr(dyh), dw=ptr(dw), N=N, H=H, W=W, C1=C1, C2=C2, Ho=H, Wo=W, R=R, S=R,
                      stride=1, pad=pad, k_slices=0, op_dtype=1)
    lib.call("b200_conv_wgrad", a, stream())
    return dw.permute(0, 3, 1, 2).cpu()


# (N, H, Cin, R, pad): ResBlock layer shapes at the three levels + ragged tails + a tile count above one wave
CASES = [(8, 28, 128, 3, 1), (3, 7, 128, 3, 1), (4, 14, 256, 3, 1), (2, 28, 256, 1, 0), (1, 7, 256, 3, 1),
         (2, 32, 128, 3, 1), (48, 28, 128, 3, 1)]


@pytest.mark.parametrize("N,H,Cin,R,pad", CASES)
def test_bf16_fprop(lib, N, H, Cin, R, pad):
    g = torch.Generator().manual_seed(N * 100 + H)
    x = bf(rn(g, N, 128, H, H))
    x2 = bf(rn(g, N, 128, H, H)) if Cin == 256 else None
    w = bf(rn(g, 128, Cin, R, R) / (Cin * R * R) ** 0.5)
    xin = x if x2 is None else torch.cat([x, x2], 1)
    ref = F.conv2d(xin, w, padding=pad)
    assert rel(fprop_bf16(lib, x, w, pad, x2=x2), ref) < TOL


def test_bf16_fprop_fused_epilogue_and_tap_order(lib):
    g = torch.Generator().manual_seed(5)
    x, w = bf(rn(g, 4, 128, 14, 14)), bf(rn(g, 128, 128, 3, 3) * 0.03)
    bias, emb, res = rn(g, 128), rn(g, 4, 128), rn(g, 4, 128, 14, 14)
    ref = F.conv2d(x, w, bias=bias, padding=1) + emb[:, :, None, None] + res
    assert rel(fprop_bf16(lib, x, w, 1, bias=bias, emb=emb, residual=res), ref) < TOL
    perm = torch.randperm(128, generator=g)
    for kh, kw in [(0, 0), (1, 1), (2, 2), (0, 2), (1, 0)]:
        w1 = torch.zeros(128, 128, 3, 3)
        w1[torch.arange(128), perm, kh, kw] = 1.0
        assert rel(fprop_bf16(lib, x, w1, 1), F.conv2d(x, w1, padding=1)) < 1e-6


@pytest.mark.parametrize("N,H,Cin,R,pad", CASES)
def test_bf16_dgrad(lib, N, H, Cin, R, pad):
    g = torch.Generator().manual_seed(N + H)
    dy = bf(rn(g, N, 128, H, H))
    w = bf(rn(g, 128, Cin, R, R) / (128 * R * R) ** 0.5)
    ref = torch.nn.grad.conv2d_input((N, Cin, H, H), w, dy, padding=pad)
    assert rel(dgrad_bf16(lib, dy, w, pad), ref) < TOL


@pytest.mark.parametrize("N,H,Cin,R,pad", CASES)
def test_bf16_wgrad(lib, N, H, Cin, R, pad):
    g = torch.Generator().manual_seed(N * 7 + H)
    x = bf(rn(g, N, 128, H, H))
    x2 = bf(rn(g, N, 128, H, H)) if Cin == 256 else None
    dy = bf(rn(g, N, 128, H, H))
    xin = x if x2 is None else torch.cat([x, x2], 1)
    ref = torch.nn.grad.conv2d_weight(xin, (128, Cin, R, R), dy, padding=pad)
    assert rel(wgrad_bf16(lib, x, dy, R, pad, x2=x2), ref) < TOL


def test_cast_bf16_strided(lib):
    g = torch.Generator().manual_seed(1)
    src = rn(g, 300, 256).to(DEV)
    dst = torch.zeros(300, 128, dtype=torch.bfloat16, device=DEV)
    a = lib.CastArgs(src=ptr(src, 128), dst=ptr(dst), rows=300, cols=128, src_ld=256)
    lib.call("b200_cast_bf16", a, stream())
    assert torch.equal(dst, src[:, 128:].to(torch.bfloat16))


@pytest.mark.parametrize("N,HW,C2", [(3, 784, 0), (2, 196, 128), (2, 4096, 0)])
def test_gn_bf16_outputs(lib, N, HW, C2):
    """y_dtype = 1 writes bf16(y); dx_bf16 is the bf16 copy of the first 128 channels of dx (fused and streaming paths)."""
    g = torch.Generator().manual_seed(HW + C2)
    Cc = 128 + C2
    x = (rn(g, N, HW, Cc) * 2 + 0.5).to(DEV)
    x1, x2 = x[..., :128].contiguous(), (x[..., 128:].contiguous() if C2 else None)
    gd, bd = (1 + 0.2 * rn(g, Cc)).to(DEV), (0.1 * rn(g, Cc)).to(DEV)
    stats = torch.zeros(N, 64, device=DEV)
    y32 = torch.empty(N, HW, Cc, device=DEV)
    y16 = torch.zeros(N, HW, Cc, dtype=torch.bfloat16, device=DEV)
    a = lib.GnArgs(x=ptr(x1), x2=ptr(x2), stats=ptr(stats), gamma=ptr(gd), beta=ptr(bd), y=ptr(y32), N=N, HW=HW, C1=128,
                   C2=C2, x_ld=0, x2_ld=0, eps=1e-5, silu=1, round_tf32=0, y_dtype=0)
    lib.call("b200_gn_fwd", a, stream())
    a.y, a.y_dtype = ptr(y16), 1
    lib.call("b200_gn_fwd", a, stream())
    # same arithmetic, different rounding of the store (+ atomics-order jitter of mean/rstd: allow 1 bf16 ulp)
    assert ((y16.float() - y32).abs() <= y32.abs() * 2 ** -8 + 1e-6).all()
    assert (y16 == y32.to(torch.bfloat16)).float().mean() > 0.999
    dy = rn(g, N, HW, Cc).to(DEV)
    dx = torch.empty(N, HW, Cc, device=DEV)
    dxb = torch.zeros(N, HW, 128, dtype=torch.bfloat16, device=DEV)
    dgam, dbet = torch.zeros(Cc, device=DEV), torch.zeros(Cc, device=DEV)
    ws = torch.zeros(N * 64, device=DEV)
    b = lib.GnBwdArgs(fwd=a, dy=ptr(dy), dx=ptr(dx), add1=None, add2=None, add1_ld=0, add2_ld=0, add_c=0,
                      dgamma=ptr(dgam), dbeta=ptr(dbet), round_dx=0, dx_bf16=ptr(dxb), dx_bf16_c=128)
    lib.call("b200_gn_bwd", b, stream(), ws=ptr(ws), ws_bytes=ws.numel() * 4)
    assert torch.equal(dxb, dx[..., :128].to(torch.bfloat16))


# ------------------------------------------------------------------------------------------------ whole model, bf16 mode
# precision='bf16': every stride-1 convolution takes bf16 operands (8 mantissa bits instead of TF32's 10) and the tensors
# only GroupNorm reads (conv1 outputs, dgrad outputs) are STORED as bf16, like the reference's autocast step.  The stated
# bars are ~8x the TF32 ones.  Measured on B200 (tools/bf16_check.py): raw net output vs the reference fixture 4.7e-3
# rms-relative / 5.5e-3 .. 6.3e-3 max-norm (TF32: 7.5e-4 / 1.0e-3); parameter gradients of a linear loss worst 2.1e-2
# max-norm, median 1.2e-2 (TF32: 3.3e-3 / 1.6e-3).  With fp32 storage (B200_BF16_STORE=0) the figures are 2.9e-3 /
# 3.3e-3 and 1.35e-2 / 7.2e-3.
BF16_NET_TOL = 8e-3       # rms-relative
BF16_NET_TOL_MAX = 1e-2   # max-norm relative
BF16_GRAD_TOL = 4e-2      # max-norm relative, worst of the 156 tensors


def rms_rel(a, b):
    return ((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt()).item()


def make_model(**over):
    from generative_models_b200.model import DiffusionModelB200, default_G
    from oracle import ddpm_oracle as O
    m = DiffusionModelB200(default_G(device=DEV, **over))
    sd = O.random_state_dict(1)
    m.net.load_state_dict(sd)
    return m, sd


def test_bf16_unet_forward_vs_reference_fixture(built_lib, golden):
    from generative_models_b200.plan import UNetPlan
    m, _ = make_model(mean_type="eps", precision="bf16")
    p = UNetPlan(m.store, 3, 28, train=False, bf16=True)
    p.z_in.copy_(golden["unet_z"].reshape(3, -1))
    p.t.copy_(golden["unet_logsnr"])
    p.guide.copy_(golden["unet_guide"])
    p.forward(m._stream())
    out = p.out.cpu().view(3, 1, 28, 28)
    assert rms_rel(out, golden["unet_out"]) < BF16_NET_TOL
    assert rel(out, golden["unet_out"]) < BF16_NET_TOL_MAX
    assert rms_rel(out, golden["unet_out"]) > 1e-4  # it really is the bf16 path (TF32 sits at 7.5e-4, fp32 at 1e-6)


def test_bf16_all_grads_vs_oracle_autograd(built_lib):
    from oracle import ddpm_oracle as O
    m, sd0 = make_model(mean_type="eps", precision="bf16")
    g = torch.Generator().manual_seed(77)
    B = 6
    x = torch.rand(B, 1, 28, 28, generator=g) * 2 - 1
    y = torch.tensor([0, 5, -1, 9, 3, -1])
    eps = torch.randn(B, 1, 28, 28, generator=g)
    u = torch.rand(B, generator=g) * 0.8 + 0.1
    r = torch.randn(B, 1, 28, 28, generator=g)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    _, _, aux = O.training_loss(sd, x, y, eps, u, "eps")
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    (O.unet_forward(sd, aux["z"], aux["logsnr"], guide=y) * r).sum().backward()
    m.store.grad.zero_()
    _, b = m.loss_and_grads(x.to(DEV), y.to(DEV), eps=eps.to(DEV), u=u.to(DEV))
    assert rms_rel(b.plan.out.cpu().view(B, 1, 28, 28), aux["out"].detach()) < BF16_NET_TOL
    m.store.grad.zero_()
    st = m._stream()
    b.plan.forward(st)
    b.plan.dout.copy_(r.reshape(B, -1))
    b.plan.backward(st)
    worst = ("", 0.0)
    for k, p in m.net.named_parameters():
        if sd[k].grad is None:
            assert p.grad.abs().max().item() == 0.0, k
            continue
        e = rel(p.grad.cpu(), sd[k].grad)
        if e > worst[1]:
            worst = (k, e)
    assert worst[1] < BF16_GRAD_TOL, worst


def test_bf16_training_tracks_tf32(built_lib):
    """same seeds, data and draws in both precisions: the loss curves agree step by step and fall."""
    from generative_models_b200.model import DiffusionModelB200, default_G
    g = torch.Generator().manual_seed(3)
    x = (torch.rand(16, 1, 28, 28, generator=g) * 2 - 1).to(DEV)
    y = torch.randint(0, 10, (16,), generator=g).to(DEV)
    curves = {}
    for prec in ("tf32", "bf16"):
        torch.manual_seed(0)
        m = DiffusionModelB200(default_G(device=DEV, mean_type="eps", precision=prec, lr=1e-3))
        torch.manual_seed(1)
        curves[prec] = [m.train_step(x, y.clone())["loss"].item() for _ in range(12)]
        assert curves[prec][-1] < curves[prec][0]
    for k in range(3):
        assert abs(curves["bf16"][k] - curves["tf32"][k]) < 1e-2 * abs(curves["tf32"][k]), (k, curves)
    assert abs(curves["bf16"][-1] - curves["tf32"][-1]) < 0.2 * abs(curves["tf32"][-1]), curves


def test_bf16_sampler_step_vs_reference_fixture(built_lib, golden):
    """teacher-forced ancestral steps (the reference's own z_t and noise): posterior sample within the bf16 pixel budget."""
    f = golden["samp_noisy_plain"]
    m, _ = make_model(mean_type=f["mean_type"], sampler=f["sampler"], timesteps=f["T"], precision="bf16")
    T, n = f["T"], f["init"].shape[0]
    noises = [t.to(DEV) for t in f["noises"]]
    b = m._plan(n, "sample")
    for k in range(T):
        z_t = f["init"] if k == 0 else f["zs"][k - 1]
        z_ours = b.teacher_step(z_t.to(DEV), f["y"].to(DEV), None, T - 1 - k, noises[k])[2].cpu()
        far = ((z_ours.double() - f["zs"][k].double()).abs() > 2e-2).double().mean().item()
        assert far < 0.03, (k, far)
        if k > 0:
            assert rel(z_ours, f["zs"][k]) < 2e-2, k


@pytest.mark.parametrize("N,HW,C2", [(3, 784, 0), (2, 196, 128), (4, 49, 0)])
def test_gn_bf16_inputs(lib, N, HW, C2):
    """x_dtype / dy_dtype = 1: bf16 sources give exactly what the fp32 kernels give on the same (bf16-representable)
    values; a NULL dx writes only the bf16 copy and the fused column sums."""
    g = torch.Generator().manual_seed(HW + C2 + 1)
    Cc = 128 + C2
    x = bf(rn(g, N, HW, Cc) * 2 + 0.5).to(DEV)
    dy = bf(rn(g, N, HW, Cc)).to(DEV)
    x1, x2 = x[..., :128].contiguous(), (x[..., 128:].contiguous() if C2 else None)
    x1b, x2b = x1.bfloat16(), (x2.bfloat16() if C2 else None)
    dyb = dy.bfloat16()
    gd, bd = (1 + 0.2 * rn(g, Cc)).to(DEV), (0.1 * rn(g, Cc)).to(DEV)
    res = {}
    for mode in (0, 1):
        stats = torch.zeros(N, 64, device=DEV)
        y = torch.empty(N, HW, Cc, device=DEV)
        a = lib.GnArgs(x=ptr(x1b if mode else x1), x2=ptr(x2b if mode else x2), stats=ptr(stats), gamma=ptr(gd),
                       beta=ptr(bd), y=ptr(y), N=N, HW=HW, C1=128, C2=C2, x_ld=0, x2_ld=0, eps=1e-5, silu=1, round_tf32=0,
                       y_dtype=0, x_dtype=mode)
        lib.call("b200_gn_fwd", a, stream())
        dx = torch.empty(N, HW, Cc, device=DEV)
        dxb = torch.zeros(N, HW, 128, dtype=torch.bfloat16, device=DEV)
        dgam, dbet = torch.zeros(Cc, device=DEV), torch.zeros(Cc, device=DEV)
        tot, per = torch.zeros(128, device=DEV), torch.zeros(N, 128, device=DEV)
        b = lib.GnBwdArgs(fwd=a, dy=ptr(dyb if mode else dy), dx=ptr(dx) if mode == 0 else None, add1=None, add2=None,
                          add1_ld=0, add2_ld=0, add_c=0, dgamma=ptr(dgam), dbeta=ptr(dbet), round_dx=0, dx_bf16=ptr(dxb),
                          dx_bf16_c=128, colsum_per_sample=ptr(per), colsum_ps_ld=128, colsum_total1=ptr(tot),
                          colsum_total2=None, colsum_c=128, dy_dtype=mode)
        lib.call("b200_gn_bwd", b, stream())
        res[mode] = (y.cpu(), stats.cpu(), dxb.float().cpu(), dgam.cpu(), dbet.cpu(), tot.cpu(), per.cpu(), dx.cpu())
    for i, (u, v) in enumerate(zip(res[0][:7], res[1][:7])):
        # same values, same arithmetic: only the atomics' order differs (the bf16 copy may flip one ulp on a few elements)
        assert rel(v, u) < (2 ** -7 if i == 2 else 2e-5), i
    # fused column sums against the fp32 dx of the fp32 run
    dx0 = res[0][7][..., :128]
    assert rel(res[0][5], dx0.sum((0, 1))) < 1e-4 and rel(res[0][6], dx0.sum(1)) < 1e-4


def test_bf16_conv_bf16_output(lib):
    """out_dtype = 1: the strip epilogue stores bf16 (fprop with emb, and dgrad into a 256-wide buffer at n_off)."""
    g = torch.Generator().manual_seed(9)
    N, H = 4, 14
    x, w = bf(rn(g, N, 128, H, H)), bf(rn(g, 128, 128, 3, 3) * 0.03)
    bias, emb = rn(g, 128), rn(g, N, 128)
    ref = F.conv2d(x, w, bias=bias, padding=1) + emb[:, :, None, None]
    xh, wk = dev_bf(nhwc(x)), dev_bf(w_khwc(w))
    out = torch.zeros(N, H, H, 128, dtype=torch.bfloat16, device=DEV)
    bd, ed = bias.to(DEV), emb.to(DEV)
    a = lib.ConvArgs(x=ptr(xh), x2=None, w=ptr(wk), out=ptr(out), bias=ptr(bd), emb=ptr(ed), residual=None, N=N, H=H, W=H,
                     C1=128, C2=0, Ho=H, Wo=H, R=3, S=3, stride=1, pad=1, mode=0, w_cout=128, w_cin=128, n_off=0, out_ld=128,
                     emb_ld=128, res_ld=128, op_dtype=1, out_dtype=1)
    lib.call("b200_conv_igemm", a, stream())
    got = out.float().permute(0, 3, 1, 2).cpu()
    assert rel(got, ref) < 2 ** -8
    w2 = bf(rn(g, 128, 256, 3, 3) * 0.03)
    dy = bf(rn(g, N, 128, H, H))
    ref2 = torch.nn.grad.conv2d_input((N, 256, H, H), w2, dy, padding=1)
    dyh, w2k = dev_bf(nhwc(dy)), dev_bf(w_khwc(w2))
    out2 = torch.zeros(N, H, H, 256, dtype=torch.bfloat16, device=DEV)
    import ctypes as C
    for n_off in (0, 128):
        a = lib.ConvArgs(x=ptr(dyh), x2=None, w=ptr(w2k), out=C.c_void_p(out2.data_ptr() + 2 * n_off), bias=None, emb=None,
                         residual=None, N=N, H=H, W=H, C1=128, C2=0, Ho=H, Wo=H, R=3, S=3, stride=1, pad=1, mode=1,
                         w_cout=128, w_cin=256, n_off=n_off, out_ld=256, emb_ld=0, res_ld=0, op_dtype=1, out_dtype=1)
        lib.call("b200_conv_igemm", a, stream())
    assert rel(out2.float().permute(0, 3, 1, 2).cpu(), ref2) < 2 ** -8


@pytest.mark.parametrize("N,H,stride", [(3, 28, 2), (2, 14, 2), (2, 7, 1)])
def test_im2col_conv_bf16_copy(lib, N, H, stride):
    """out_bf16 on the layers the im2col kernel takes (the stride-2 Downsample convs, small stride-1 levels): the dense
    bf16 copy equals the fp32 output rounded to bf16."""
    g = torch.Generator().manual_seed(H + stride)
    x = rn(g, N, 128, H, H)
    w = rn(g, 128, 128, 3, 3) * 0.03
    Ho = (H + 2 - 3) // stride + 1
    xh, wk = nhwc(x).to(DEV), w_khwc(w).to(DEV)
    out = torch.full((N, Ho, Ho, 128), float("nan"), device=DEV)
    outb = torch.zeros(N, Ho, Ho, 128, dtype=torch.bfloat16, device=DEV)
    a = lib.ConvArgs(x=ptr(xh), x2=None, w=ptr(wk), out=ptr(out), bias=None, emb=None, residual=None, N=N, H=H, W=H,
                     C1=128, C2=0, Ho=Ho, Wo=Ho, R=3, S=3, stride=stride, pad=1, mode=0, w_cout=128, w_cin=128, n_off=0,
                     out_ld=128, emb_ld=0, res_ld=0, out_bf16=ptr(outb))
    lib.call("b200_conv_igemm", a, stream())
    ref = F.conv2d(x, w, stride=stride, padding=1)
    assert rel(out.permute(0, 3, 1, 2).cpu(), ref) < 2e-3
    assert torch.equal(outb, out.to(torch.bfloat16))
