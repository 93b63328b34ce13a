"""-m gpu parity of the bandwidth kernels (through the C ABI) against the CPU oracle on the same seeded inputs.
Tolerances: q_sample and the sampler update are a handful of fp32 ops per pixel and are held to 1e-5 relative
(device expf/sqrt vs CPU libm differ in the last ulp, so bit-exactness vs a CPU oracle is not defined); the loss sums
784 terms (1e-5); gradients 1e-4."""
import pytest
import torch

from oracle import ddpm_oracle as O
from tests.gpu_util import ptr, rel, stream

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def lib(built_lib):
    built_lib.load().b200_init(0)
    return built_lib


def _inputs(B, D, seed, lo=-15.0, hi=15.0):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, D, generator=g) * 2 - 1
    eps = torch.randn(B, D, generator=g)
    logsnr = torch.rand(B, generator=g) * (hi - lo) + lo
    return x, eps, logsnr


@pytest.mark.parametrize("B,D", [(1, 784), (7, 784), (64, 1024), (513, 784)])
def test_q_sample(lib, B, D):
    x, eps, l = _inputs(B, D, 1, -20.0, 20.0)
    ref = O.q_sample(x, l, eps)
    xd, ed, ld = x.to(DEV), eps.to(DEV), l.to(DEV)
    z = torch.full((B, D), float("nan"), device=DEV)
    lib.call("b200_q_sample", lib.QSampleArgs(x=ptr(xd), eps=ptr(ed), logsnr=ptr(ld), z=ptr(z), B=B, D=D), stream())
    assert rel(z.cpu(), ref) < 1e-5
    # against the same torch expression on the device (the reference's GPU path): a few ulp (FMA contraction aside)
    ref_d = O.q_sample(xd, ld, ed)
    assert rel(z, ref_d) < 1e-6


def test_q_sample_rejects_bad_args(lib):
    a = lib.QSampleArgs(x=None, eps=None, logsnr=None, z=None, B=1, D=784)
    with pytest.raises(lib.B200Error):
        lib.call("b200_q_sample", a, stream())


@pytest.mark.parametrize("mean_type", ["eps", "x", "v"])
@pytest.mark.parametrize("B", [3, 64])
def test_loss_fwd_bwd(lib, mean_type, B):
    D = 784
    x, eps, l = _inputs(B, D, 2, -8.0, 8.0)
    g = torch.Generator().manual_seed(3)
    z = O.q_sample(x, l, eps)
    # a plausible net output: the true target plus noise, so both clip branches and both max branches occur
    tgt = {"eps": eps, "x": x, "v": torch.sqrt(torch.sigmoid(l))[:, None] * eps - torch.sqrt(torch.sigmoid(-l))[:, None] * x}[mean_type]
    out = (tgt + 0.3 * torch.randn(B, D, generator=g)).requires_grad_(True)
    per = O.loss_from_out(out, z, x, eps, l, mean_type)
    per.mean().backward()
    d = [t.to(DEV) for t in (out.detach(), z, x, eps, l)]
    loss_b = torch.zeros(B, device=DEV)
    mse2 = torch.zeros(B, 2, device=DEV)
    dout = torch.full((B, D), float("nan"), device=DEV)
    mean = torch.zeros(1, device=DEV)
    a = lib.LossArgs(out=ptr(d[0]), z=ptr(d[1]), x=ptr(d[2]), eps=ptr(d[3]), logsnr=ptr(d[4]), loss_b=ptr(loss_b),
                     mse2=ptr(mse2), dout=ptr(dout), loss_mean=ptr(mean), B=B, D=D,
                     mean_type={"eps": 0, "x": 1, "v": 2}[mean_type], grad_scale=1.0 / B)
    lib.call("b200_loss", a, stream())
    assert rel(loss_b.cpu(), per.detach()) < 1e-5
    assert abs(mean.item() - per.mean().item()) < 1e-5 * abs(per.mean().item())
    assert rel(dout.cpu(), out.grad) < 1e-4


def _sampler_case(lib, mean_type, sampler, guided, step, T=6, n=5, D=784):
    from generative_models_b200.schedule import sampler_table
    g = torch.Generator().manual_seed(4 + step)
    z = torch.randn(n, D, generator=g)
    out_c = torch.randn(n, D, generator=g)
    out_u = torch.randn(n, D, generator=g)
    noise = torch.randn(n, D, generator=g)
    w = 4 * torch.rand(n, generator=g)
    sched = sampler_table(T, "cpu")
    lt, ls = sched[step, 0], sched[step, 1]
    ltv = lt.expand(n)
    xc, ec = O.run_model(out_c, z, ltv, mean_type)
    if guided:
        _, eu = O.run_model(out_u, z, ltv, mean_type)
        xc, ec = O.cf_guidance(ec, eu, z, ltv, w)
    if sampler == "noisy":
        zs, _, _ = O.posterior_noisy(z, xc, lt, ls, noise)
    else:
        zs = O.posterior_ddim(xc, ec, ls)
    if step == 0:
        zs = xc
    dz, dc, du, dn, dw, dsch = [t.to(DEV) for t in (z, out_c, out_u, noise, w, sched)]
    zn = torch.full((n, D), float("nan"), device=DEV)
    xh = torch.zeros(n, D, device=DEV)
    eh = torch.zeros(n, D, device=DEV)
    step_dev = torch.tensor([step], dtype=torch.int32, device=DEV)
    a = lib.SamplerStepArgs(z=ptr(dz), out_c=ptr(dc), out_u=ptr(du) if guided else None, cond_w=ptr(dw) if guided else None,
                            noise=ptr(dn), sched=ptr(dsch), step_dev=ptr(step_dev), z_next=ptr(zn), x_hist=ptr(xh),
                            e_hist=ptr(eh), n=n, D=D, step=-1, mean_type={"eps": 0, "x": 1, "v": 2}[mean_type],
                            sampler={"ddim": 0, "noisy": 1}[sampler])
    lib.call("b200_sampler_step", a, stream())
    return (zn.cpu(), zs), (xh.cpu(), xc), (eh.cpu(), ec)


@pytest.mark.parametrize("mean_type", ["eps", "v", "x"])
@pytest.mark.parametrize("sampler", ["noisy", "ddim"])
@pytest.mark.parametrize("guided", [False, True])
@pytest.mark.parametrize("step", [0, 2, 5])
def test_sampler_step(lib, mean_type, sampler, guided, step):
    (zn, zs), (xh, xc), (eh, ec) = _sampler_case(lib, mean_type, sampler, guided, step)
    assert rel(xh, xc) < 1e-5          # posterior-mean input x~ (clipped): well conditioned
    assert rel(zn, zs) < 2e-5          # posterior mean + std*noise (or the ddim update)
    # eps re-derived from the clipped x is ill-conditioned at logsnr -> +20: abs error scaled by its range
    assert rel(eh, ec) < 2e-3


def test_sampler_step_histories_and_fused_proc(lib):
    """hist_T > 0: the step writes row T-1-index of the [T][n*D] stacks (gaussian_diffusion.py:293-296); the uint8 stacks
    hold diffusion_model.py:91-92 proc() of the same values bit for bit; z_dup mirrors z_next (2n-row guided batch)."""
    from generative_models_b200.schedule import sampler_table
    g = torch.Generator().manual_seed(3)
    n, D, T, step = 3, 784, 6, 2
    z, out = torch.randn(n, D, generator=g).to(DEV), torch.randn(n, D, generator=g).to(DEV)
    noise = torch.randn(n, D, generator=g).to(DEV)
    sched = sampler_table(T, DEV)
    zn, zd = torch.zeros(n, D, device=DEV), torch.zeros(n, D, device=DEV)
    f32 = [torch.full((T, n, D), 7.0, device=DEV) for _ in range(3)]
    u8 = [torch.full((T, n, D), 9, dtype=torch.uint8, device=DEV) for _ in range(3)]
    a = lib.SamplerStepArgs(z=ptr(z), out_c=ptr(out), out_u=None, cond_w=None, noise=ptr(noise), sched=ptr(sched),
                            step_dev=None, z_next=ptr(zn), z_dup=ptr(zd), z_hist=ptr(f32[0]), x_hist=ptr(f32[1]),
                            e_hist=ptr(f32[2]), hist_T=T, z_u8=ptr(u8[0]), x_u8=ptr(u8[1]), e_u8=ptr(u8[2]), n=n, D=D,
                            step=step, mean_type=0, sampler=1)
    lib.call("b200_sampler_step", a, stream())
    row = T - 1 - step
    assert torch.equal(zd, zn) and torch.equal(f32[0][row], zn)
    for h32, h8 in zip(f32, u8):
        want = ((h32[row] + 1) * 127.5).clamp(0, 255).to(torch.uint8)
        assert torch.equal(h8[row], want)
        for r in range(T):
            if r != row:
                assert (h32[r] == 7.0).all() and (h8[r] == 9).all()
    assert f32[1][row].abs().max() <= 1.0  # the clipped x prediction
    a.step = T  # outside the history
    assert lib.load().b200_sampler_step(__import__("ctypes").byref(a), None, 0, stream()) < 0


def test_adam_reads_device_hparams(lib):
    """hparams_dev = {lr, grad_scale} in device memory overrides the struct fields (a captured step graph follows an
    lr schedule without re-capture)."""
    g = torch.Generator().manual_seed(0)
    n = 4096
    p0, gr = torch.randn(n, generator=g), torch.randn(n, generator=g)
    outs = []
    for use_dev in (False, True):
        p, m, v = p0.clone().to(DEV), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
        hp = torch.tensor([1e-2, 0.5], device=DEV)
        a = lib.AdamArgs(p=ptr(p), g=ptr(gr.to(DEV)), m=ptr(m), v=ptr(v), n=n, lr=1e-2 if not use_dev else 123.0,
                         beta1=0.9, beta2=0.999, eps=1e-8, grad_scale=0.5 if not use_dev else 77.0, step_dev=None, step=1,
                         hparams_dev=ptr(hp) if use_dev else None)
        lib.call("b200_adam", a, stream())
        outs.append(p.cpu())
    assert torch.equal(outs[0], outs[1])
    assert (outs[0] - p0).abs().max() > 5e-3


def test_step_advance_and_increment(lib):
    v = torch.tensor([5], dtype=torch.int32, device=DEV)
    lib.call("b200_step_advance", lib.StepAdvanceArgs(step_dev=ptr(v)), stream())
    assert v.item() == 4
    lib.call("b200_int_increment", lib.StepAdvanceArgs(step_dev=ptr(v)), stream())
    lib.call("b200_int_increment", lib.StepAdvanceArgs(step_dev=ptr(v)), stream())
    assert v.item() == 6


def test_adam_matches_torch(lib):
    n = 100_003
    g = torch.Generator().manual_seed(9)
    p0 = torch.randn(n, generator=g)
    ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=3e-4)
    p = p0.to(DEV)
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    step = torch.zeros(1, dtype=torch.int32, device=DEV)
    for it in range(3):
        gr = torch.randn(n, generator=g) * (10.0 ** (it - 1))
        ref.grad = gr.clone()
        opt.step()
        gd = (gr * 2).to(DEV)  # grad_scale 0.5 undoes the factor 2 (the 1/world_size path)
        lib.call("b200_int_increment", lib.StepAdvanceArgs(step_dev=ptr(step)), stream())
        a = lib.AdamArgs(p=ptr(p), g=ptr(gd), m=ptr(m), v=ptr(v), n=n, lr=3e-4, beta1=0.9, beta2=0.999, eps=1e-8,
                         grad_scale=0.5, step_dev=ptr(step), step=0)
        lib.call("b200_adam", a, stream())
        assert (p.cpu() - ref.detach()).abs().max().item() < 2e-6


@pytest.mark.parametrize("M,N,K", [(6, 128, 512), (3, 512, 128), (512, 1664, 512), (500, 512, 128), (512, 512, 512)])
@pytest.mark.parametrize("act", [0, 1])
def test_linear_fwd_bwd_vs_torch(lib, M, N, K, act):
    """b200_linear (the embedding MLPs: simple_unet.py:33-37 time_embed, :119-121 emb_layers) in its three modes against
    torch fp32 on the CPU: Y = silu?(X) W^T + b, dX = dY W (* silu'(X)), dW += dY^T silu?(X), db += colsum(dY).  The
    batch-512 shapes take the 64x64-tile kernel (500 = a ragged last tile), the small ones the 16x16 one.  fp32 FMAs
    in a different summation order: 1e-5 relative."""
    g = torch.Generator().manual_seed(M * 7 + N + act)
    X = torch.randn(M, K, generator=g)
    W = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    dY = torch.randn(M, N, generator=g)
    Xr = X.clone().requires_grad_(True)
    Wr = W.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    Yr = torch.nn.functional.linear(torch.nn.functional.silu(Xr) if act else Xr, Wr, br)
    Yr.backward(dY)
    d = {k: v.to(DEV) for k, v in dict(X=X, W=W, b=b, dY=dY).items()}
    Y = torch.empty(M, N, device=DEV)
    dX = torch.empty(M, K, device=DEV)
    dW = torch.zeros(N, K, device=DEV)
    db = torch.zeros(N, device=DEV)
    base = dict(X=ptr(d["X"]), W=ptr(d["W"]), bias=ptr(d["b"]), Y=ptr(Y), dY=ptr(d["dY"]), dX=ptr(dX), dW=ptr(dW),
                db=ptr(db), M=M, N=N, K=K, act=act)
    for mode in (0, 1, 2):
        lib.call("b200_linear", lib.LinearArgs(mode=mode, **base), stream())
    torch.cuda.synchronize()
    assert rel(Y.cpu(), Yr.detach()) < 1e-5
    assert rel(dX.cpu(), Xr.grad) < 1e-5
    assert rel(dW.cpu(), Wr.grad) < 1e-5
    assert rel(db.cpu(), br.grad) < 1e-5
