"""-m gpu: progressive distillation (gaussian_diffusion.py:105-154, simple_unet.py:60-64,
diffusion_model.py:34-43) against fixtures produced by the reference's own DiffusionModel with a teacher
(oracle/make_golden.py §7) and against the pinned oracle.

  * step1 / step2 distillation losses and gradients (all 160 tensors now live: the four cond_w_embed ones included);
  * the cond_w embedding path alone (linear loss: the tight check of its forward and backward);
  * the teacher-target kernels (per-sample log-SNR DDIM step, b200_distill_target) against the oracle on the same inputs;
  * sampling of a distilled student (conditioned on w, unguided) and sampler='teacher_test'.
"""
import pytest
import torch

from oracle import ddpm_oracle as O
from tests.gpu_util import rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
NET_TOL = 1e-3          # rms-relative raw net output (TF32), as in tests/test_gpu_model.py
GRAD_TOL = 6e-3         # max-norm, gradients of a linear loss
LOSS_GRAD_TOL = 1e-1    # gradients of the clipped diffusion loss (clip mask flips)


def rms_rel(a, b):
    return ((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt()).item()


def make_distill_model(tmp_path, mode, mean_type, T, sampler="ddim", **over):
    from generative_models_b200.model import DiffusionModelB200, default_G
    tp = tmp_path / "teacher.pt"
    torch.save({"net." + k: v for k, v in O.random_state_dict(2).items()}, tp)
    G = default_G(device=DEV, mean_type=mean_type, sampler=sampler, timesteps=T, teacher_path=tp, teacher_mode=mode, **over)
    m = DiffusionModelB200(G)
    assert m.teacher_store is not None
    # the student starts as a copy of the teacher (diffusion_model.py:36-40) ...
    assert torch.equal(m.store.flat, m.teacher_store.flat)
    sd_t = O.random_state_dict(2)
    assert torch.equal(m.state_dict()["teacher_net.turn.in_layers.2.weight"].cpu(), sd_t["turn.in_layers.2.weight"])
    assert len(m.state_dict()) == 320 and not any(p.requires_grad for p in m.teacher_net.parameters())
    # ... the fixtures then give it different weights so that the targets are not met trivially
    m.net.load_state_dict(O.random_state_dict(1))
    return m


@pytest.mark.parametrize("tag", ["distill_step1_eps", "distill_step2_eps", "distill_step2_v"])
def test_distillation_loss_and_grads_vs_reference_fixture(built_lib, golden, tmp_path, tag):
    f = golden[tag]
    m = make_distill_model(tmp_path, f["mode"], f["mean_type"], f["T"])
    dev = lambda t: None if t is None else t.to(DEV)
    kw = dict(eps=dev(f["eps"]), cond_w=dev(f["cond_w"]))
    kw.update(idx=dev(f["i"])) if f["mode"] == "step2" else kw.update(u=dev(f["u"]))
    loss, _ = m.loss(dev(f["x"]), dev(f["y"]), **kw)
    ref = f["loss"].item()
    assert abs(loss.item() - ref) < 5e-3 * abs(ref), (loss.item(), ref)
    m.store.grad.zero_()
    loss2, b = m.loss_and_grads(dev(f["x"]), dev(f["y"]), **kw)
    assert abs(loss2.item() - ref) < 5e-3 * abs(ref)
    # the teacher's targets against the oracle's (same teacher weights, z_t, logsnr, cond_w)
    sd_t = O.random_state_dict(2)
    u = (f["i"] + 1).float() / f["T"] if f["mode"] == "step2" else f["u"]
    logsnr = O.logsnr_cosine(u)
    z = O.q_sample(f["x"], logsnr, f["eps"])
    x_t, e_t = O.distill_targets(sd_t, z, f["y"], logsnr, u, f["cond_w"], f["mode"], f["T"], f["mean_type"], i=f["i"])
    n = f["x"].shape[0]
    far = lambda a, b_: ((a - b_).abs() > 5e-3).double().mean().item()
    assert far(b.x_target.cpu().view(n, 1, 28, 28), x_t) < 0.02
    grads = {k: p.grad for k, p in m.net.named_parameters()}
    for k, nrm in f["gradnorms"].items():
        assert abs(grads[k].norm().item() - nrm.item()) < LOSS_GRAD_TOL * nrm.item() + 1e-8, (k, grads[k].norm().item(), nrm.item())
    for k, g in f["grads"].items():
        assert rms_rel(grads[k].cpu(), g) < LOSS_GRAD_TOL, k
    assert all(grads[k].abs().max().item() > 0 for k in grads if k.startswith("cond_w_embed."))
    # the teacher is frozen
    before = m.teacher_store.flat.clone()
    m.train_step(dev(f["x"]), dev(f["y"]).clone())
    m.train_step(dev(f["x"]), dev(f["y"]).clone())
    m.train_step(dev(f["x"]), dev(f["y"]).clone())  # eager, capture, replay
    assert torch.equal(m.teacher_store.flat, before)


def test_cond_w_embedding_forward_and_backward_vs_oracle(built_lib):
    """A net conditioned on w (simple_unet.py:60-64, max_period = 4): raw output and ALL 160 gradients of a linear loss."""
    from generative_models_b200.model import DiffusionModelB200, default_G
    from generative_models_b200.plan import UNetPlan
    m = DiffusionModelB200(default_G(device=DEV, mean_type="eps"))
    sd0 = O.random_state_dict(1)
    m.net.load_state_dict(sd0)
    g = torch.Generator().manual_seed(8)
    B = 5
    z = torch.randn(B, 1, 28, 28, generator=g)
    l = torch.tensor([-6.0, -1.0, 0.2, 3.0, 9.0])
    y = torch.tensor([2, -1, 7, 0, -1])
    cw = 4.0 * torch.rand(B, generator=g)
    r = torch.randn(B, 1, 28, 28, generator=g)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    out = O.unet_forward(sd, z, l, guide=y, cond_w=cw)
    (out * r).sum().backward()
    p = UNetPlan(m.store, B, 28, train=True, cond_w=True)
    p.z_in.copy_(z.reshape(B, -1))
    p.t.copy_(l)
    p.guide.copy_(y)
    p.cond_w.copy_(cw)
    st = m._stream()
    m.store.grad.zero_()
    p.forward(st)
    ours = p.out.cpu().view(B, 1, 28, 28)
    assert rms_rel(ours, out.detach()) < NET_TOL
    # without the embedding the output differs: the branch is live
    p0 = UNetPlan(m.store, B, 28, train=False)
    p0.z_in.copy_(z.reshape(B, -1)); p0.t.copy_(l); p0.guide.copy_(y)
    p0.forward(st)
    assert rms_rel(p0.out.cpu().view(B, 1, 28, 28), out.detach()) > 1e-2
    p.dout.copy_(r.reshape(B, -1))
    p.backward(st)
    worst = ("", 0.0)
    for k, prm in m.net.named_parameters():
        assert sd[k].grad is not None, k
        e = rel(prm.grad.cpu(), sd[k].grad)
        worst = max(worst, (k, e), key=lambda t: t[1])
    assert worst[1] < GRAD_TOL, worst


def test_distill_target_kernel_vs_oracle(built_lib):
    """b200_distill_target against gaussian_diffusion.py:137-154 restated in the oracle, incl. an i == 0 row."""
    import ctypes as C
    lib = built_lib
    g = torch.Generator().manual_seed(4)
    n, D, T = 6, 784, 8
    i = torch.tensor([0, 1, 3, 7, 0, 5])
    u = (i + 1).float() / T
    lt, ls = O.logsnr_cosine(u), O.logsnr_cosine(u - 1.0 / T)
    z_t, z_teacher, x_pred = (torch.randn(n, D, generator=g) for _ in range(3))
    bc = lambda v: v[:, None]
    alpha_s, alpha_t = bc(torch.sqrt(torch.sigmoid(ls))), bc(torch.sqrt(torch.sigmoid(lt)))
    f = bc(torch.exp(0.5 * (torch.nn.functional.softplus(lt) - torch.nn.functional.softplus(ls))))
    x_ref = torch.where(bc(i == 0), x_pred, (z_teacher - f * z_t) / (alpha_s - f * alpha_t))
    e_ref = O.eps_from_x(z_t, x_ref, lt)
    d = lambda t: t.to(DEV).contiguous()
    zt_d, z_d, xp_d, lt_d, ls_d, i_d = d(z_teacher), d(z_t), d(x_pred), d(lt), d(ls), d(i)
    xo, eo = torch.zeros(n, D, device=DEV), torch.zeros(n, D, device=DEV)
    p = lambda t: C.c_void_p(t.data_ptr())
    a = lib.DistillTargetArgs(z_teacher=p(zt_d), z_t=p(z_d), x_pred_teacher=p(xp_d), logsnr_t=p(lt_d), logsnr_s=p(ls_d),
                              index=p(i_d), x_target=p(xo), eps_target=p(eo), n=n, D=D)
    lib.call("b200_distill_target", a, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rel(xo.cpu(), x_ref) < 1e-4 and torch.equal(xo.cpu()[0], x_pred[0])
    assert rel(eo.cpu(), e_ref) < 2e-3


def test_distilled_student_and_teacher_test_sampling_vs_reference_fixture(built_lib, golden, tmp_path):
    f = golden["distill_step1_eps"]
    T = f["T"]
    for key, sampler in (("sample", "ddim"), ("teacher_test", "teacher_test")):
        s = f[key]
        m = make_distill_model(tmp_path, "step1", "eps", T, sampler=sampler)
        w = s["net_cond_w"].to(DEV)
        cond_w, net_cond_w = (w, None) if sampler == "teacher_test" else (None, w)
        zs, xs, es = m.sample_trajectory(s["init"].to(DEV), y=s["y"].to(DEV), cond_w=cond_w, net_cond_w=net_cond_w)
        assert zs.shape == s["zs"].shape
        far = lambda a, b_: ((a.cpu() - b_).abs() > 5e-3).double().mean().item()
        assert far(xs, s["xs"]) < 0.05 and far(zs, s["zs"]) < 0.05, (key, far(xs, s["xs"]), far(zs, s["zs"]))
        # the public call draws like the reference: randn(n,1,S,S), then 4*rand(n) (gaussian_diffusion.py:247-251)
        torch.manual_seed(5)
        out = m.sample(2, s["y"].to(DEV))
        assert out.shape == (2, 1, 28, 28) and torch.isfinite(out).all() and out.abs().max() <= 1.0
        assert any(k[0].endswith("_t" if sampler == "teacher_test" else "_w") for k in m._plans)
