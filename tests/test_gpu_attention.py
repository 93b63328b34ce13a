"""-m gpu: the self-attention EXTENSION (csrc/attn_small.cu, DiffusionModelB200 G.attention=1).

The reference's SimpleUnet has no attention (simple_unet.py:16-72), so parity here is "unpinned": the checker is
oracle.attention_block (guided-diffusion's single-head AttentionBlock) and torch.matmul / torch.softmax for the kernels.
  * b200_bgemm (tcgen05 TF32, batched) in all four operand orientations, with ragged sizes (49 tokens) — against fp64
    matmul of the same TF32-rounded operands;
  * b200_softmax_fwd / bwd against torch;
  * the UNet with the block switched on: raw output and every parameter gradient of a linear loss against oracle
    autograd, at 7x7 tokens (1x28x28) and 16x16 tokens (3x64x64, BASELINE.json configs[4]); train_step / sample.
"""
import ctypes as C
import math

import pytest
import torch

from oracle import ddpm_oracle as O
from tests.gpu_util import rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def tf32(t):
    """round to nearest TF32 (10 mantissa bits), like the producers of every tensor-core operand do"""
    i = t.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)


def p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def rms_rel(a, b):
    return ((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt()).item()


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K,batch", [(49, 49, 128, 3), (49, 128, 49, 3), (256, 256, 128, 2), (256, 128, 256, 2), (130, 40, 72, 2)])
def test_bgemm_all_orientations(built_lib, a_mn, b_mn, M, N, K, batch):
    lib = built_lib
    g = torch.Generator().manual_seed(M + N + K)
    pad = lambda n: (n + 3) // 4 * 4 + 4  # row pitches: multiples of 4 floats, deliberately wider than the extent
    A = tf32(torch.randn(batch, *( (K, pad(M)) if a_mn else (M, pad(K)) ), generator=g)).to(DEV)
    Bm = tf32(torch.randn(batch, *( (K, pad(N)) if b_mn else (N, pad(K)) ), generator=g)).to(DEV)
    ldc = pad(N)
    Cm = torch.full((batch, M, ldc), 7.0, device=DEV)
    a = lib.BgemmArgs(A=p(A), B=p(Bm), C=p(Cm), M=M, N=N, K=K, batch=batch, lda=A.shape[2], ldb=Bm.shape[2], ldc=ldc,
                      strideA=A.shape[1] * A.shape[2], strideB=Bm.shape[1] * Bm.shape[2], strideC=M * ldc, a_mn=a_mn, b_mn=b_mn,
                      alpha=0.5, round_out=0)
    lib.call("b200_bgemm", a, stream())
    Ad = (A[:, :K, :M].transpose(1, 2) if a_mn else A[:, :M, :K]).double()
    Bd = (Bm[:, :K, :N].transpose(1, 2) if b_mn else Bm[:, :N, :K]).double()
    ref = 0.5 * Ad @ Bd.transpose(1, 2)
    assert rel(Cm[:, :, :N].cpu(), ref.cpu()) < 2e-5, (a_mn, b_mn)
    assert (Cm[:, :, N:] == 7.0).all()  # nothing is stored past N


def test_softmax_fwd_bwd(built_lib):
    lib = built_lib
    g = torch.Generator().manual_seed(0)
    for T, ld, rows in ((49, 64, 3 * 49), (256, 256, 2 * 256), (5, 32, 7)):
        s = (torch.randn(rows, ld, generator=g) * 3).to(DEV)
        pr = torch.full((rows, ld), 9.0, device=DEV)
        lib.call("b200_softmax_fwd", lib.SoftmaxArgs(s=p(s), p=p(pr), dp=None, ds=None, rows=rows, T=T, ld=ld), stream())
        ref = torch.softmax(s[:, :T].double(), -1)
        assert rel(pr[:, :T].cpu(), ref.cpu()) < 6e-4  # TF32-rounded output
        assert (pr[:, T:] == 0).all()
        dp = torch.randn(rows, ld, generator=g).to(DEV)
        ds = torch.full((rows, ld), 9.0, device=DEV)
        lib.call("b200_softmax_bwd", lib.SoftmaxArgs(s=None, p=p(pr), dp=p(dp), ds=p(ds), rows=rows, T=T, ld=ld), stream())
        pd, dd = pr[:, :T].double(), dp[:, :T].double()
        ref = pd * (dd - (pd * dd).sum(-1, keepdim=True))
        assert rel(ds[:, :T].cpu(), ref.cpu()) < 6e-4
        assert (ds[:, T:] == 0).all()


@pytest.mark.parametrize("S,Ci,prec", [(28, 1, "tf32"), (28, 1, "bf16"), (64, 3, "tf32")])
def test_unet_with_attention_vs_oracle(built_lib, S, Ci, prec):
    from generative_models_b200.model import DiffusionModelB200, default_G
    from generative_models_b200.plan import UNetPlan
    # cf_drop_prob = 0: train_step's label drop draws from the global CPU RNG (diffusion_model.py:67); the oracle loss
    # below is computed for the labels as given (with the default 0.1 and three samples the test failed one run in four)
    G = default_G(device=DEV, mean_type="eps", sampler="noisy", timesteps=3, precision=prec, img_channels=Ci, image_size=S,
                  attention=1, cf_drop_prob=0.0)
    m = DiffusionModelB200(G)
    assert len(m.state_dict()) == 170 and m.state_dict()["net.attn.proj_out.weight"].abs().max().item() == 0.0  # zero-init
    sd0 = O.random_state_dict(1, img_channels=Ci, attention=True)
    m.net.load_state_dict(sd0)
    g = torch.Generator().manual_seed(S)
    B = 3
    z = torch.randn(B, Ci, S, S, generator=g)
    logsnr = torch.tensor([1.5, -3.0, 0.2])
    y = torch.tensor([7, -1, 2])
    r = torch.randn(B, Ci, S, S, generator=g)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    out_ref = O.unet_forward(sd, z, logsnr, guide=y)
    (out_ref * r).sum().backward()
    pl = UNetPlan(m.store, B, S, train=True, bf16=(prec == "bf16"))
    pl.z_in.copy_(z.reshape(B, -1)); pl.t.copy_(logsnr); pl.guide.copy_(y)
    st = m._stream()
    m.store.grad.zero_()
    pl.forward(st)
    net_tol, grad_tol = (1e-3, 6e-3) if prec == "tf32" else (8e-3, 4e-2)
    e = rms_rel(pl.out.cpu().view(B, Ci, S, S), out_ref.detach())
    assert e < net_tol, e
    pl.dout.copy_(r.reshape(B, -1))
    pl.backward(st)
    worst = ("", 0.0)
    for k, prm in m.net.named_parameters():
        if sd[k].grad is None or k == "attn.k.bias":
            continue
        worst = max(worst, (k, rel(prm.grad.cpu(), sd[k].grad)), key=lambda t: t[1])
    assert worst[1] < grad_tol, worst
    params = dict(m.net.named_parameters())
    for k in ("attn.q.weight", "attn.q.bias", "attn.k.weight", "attn.v.weight", "attn.v.bias", "attn.proj_out.weight",
              "attn.proj_out.bias", "attn.norm.weight", "attn.norm.bias"):
        assert rel(params[k].grad.cpu(), sd[k].grad) < grad_tol, k
    # the key bias has NO gradient mathematically (softmax is invariant to a constant added to every score of a row):
    # both sides hold rounding noise, so it is bounded against the query bias's gradient instead of compared relatively
    assert sd["attn.k.bias"].grad.abs().max() < 1e-5 * sd["attn.q.bias"].grad.abs().max()
    assert params["attn.k.bias"].grad.abs().max().item() < 2e-3 * params["attn.q.bias"].grad.abs().max().item()
    # the public API
    x = (torch.rand(B, Ci, S, S, generator=g) * 2 - 1).to(DEV)
    eps, u = torch.randn(B, Ci, S, S, generator=g), torch.tensor([0.3, 0.7, 0.5])
    loss = m.train_step(x, y.to(DEV).clone(), eps=eps.to(DEV), u=u.to(DEV))["loss"].item()
    ref_loss, _, _ = O.training_loss(sd0, x.cpu(), y, eps, u, "eps")
    assert abs(loss - ref_loss.item()) < (3e-3 if prec == "tf32" else 3e-2) * abs(ref_loss.item()), (loss, ref_loss.item())
    out = m.sample(2, y[:2].to(DEV))
    assert out.shape == (2, Ci, S, S) and torch.isfinite(out).all()
