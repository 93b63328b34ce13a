"""The CPU oracle (oracle/ddpm_oracle.py) against fixtures produced by the reference itself (oracle/make_golden.py).
Both sides are CPU fp32 torch, so agreement is expected to ~1e-6; tolerances are written per check."""
import torch
import pytest

from oracle import ddpm_oracle as O


def rel(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def test_schedule_bit_exact(golden):
    assert torch.equal(O.logsnr_cosine(golden["sched_u"]), golden["sched_logsnr"])
    T = 8
    rows = torch.stack([torch.stack([O.logsnr_cosine((torch.tensor(i) + 1.0) / T), O.logsnr_cosine(torch.tensor(i) / T)])
                        for i in range(T)])
    assert torch.equal(rows, golden["sched_T8"])


def test_unet_forward(golden):
    sd = O.random_state_dict(1)
    out = O.unet_forward(sd, golden["unet_z"], golden["unet_logsnr"], guide=golden["unet_guide"])
    assert rel(out, golden["unet_out"]) < 1e-5
    out = O.unet_forward(sd, golden["unet_z"], golden["unet_logsnr"])
    assert rel(out, golden["unet_out_noguide"]) < 1e-5


def test_loss_and_grads(golden):
    for mt in ("eps", "x", "v"):
        sd = {k: v.clone().requires_grad_(True) for k, v in O.random_state_dict(1).items()}
        loss, per, _ = O.training_loss(sd, golden[f"loss_{mt}_x"], golden[f"loss_{mt}_y"], golden[f"loss_{mt}_eps"],
                                       golden[f"loss_{mt}_u"], mean_type=mt)
        assert abs(loss.item() - golden[f"loss_{mt}"].item()) <= 1e-5 * abs(golden[f"loss_{mt}"].item())
        loss.backward()
        for k, n in golden[f"loss_{mt}_gradnorms"].items():
            g = sd[k].grad
            assert g is not None, k
            assert abs(g.norm().item() - n.item()) <= 2e-4 * n.item() + 1e-9, (mt, k)
        for key in golden:
            pre = f"loss_{mt}_grad_"
            if key.startswith(pre):
                assert rel(sd[key[len(pre):]].grad, golden[key]) < 2e-4, key
        # cond_w_embed never receives a gradient on this path
        assert sd["cond_w_embed.0.weight"].grad is None


def test_sampler_trajectories(golden):
    sd = O.random_state_dict(1)
    for tag in ("noisy_guided", "ddim_plain", "noisy_plain"):
        f = golden[f"samp_{tag}"]

        def net(z, l, uncond, y=f["y"]):
            with torch.no_grad():
                return O.unet_forward(sd, z, l, guide=-torch.ones_like(y) if uncond else y)

        zs, xs, es = O.sample_loop(net, f["init"], f["T"], sampler=f["sampler"], mean_type=f["mean_type"],
                                   cond_w=f["cond_w"], noises=f["noises"])
        assert rel(zs, f["zs"]) < 1e-4, tag
        assert rel(xs, f["xs"]) < 1e-4, tag
        # eps at logsnr ~ 20 is ill-conditioned: looser
        assert rel(es, f["es"]) < 1e-3, tag


def test_adam_step(golden):
    """One reference train_step == oracle loss/backward + torch.optim.Adam defaults (diffusion_model.py:56,63-74)."""
    sd0 = O.random_state_dict(1)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    loss, _, _ = O.training_loss(sd, golden["step_x"], golden["step_y"], golden["step_eps"], golden["step_u"], "eps")
    assert abs(loss.item() - golden["step_loss"].item()) <= 1e-5 * abs(golden["step_loss"].item())
    loss.backward()
    params = [p for p in sd.values() if p.grad is not None]
    torch.optim.Adam(params, lr=3e-4).step()
    for key in golden:
        if key.startswith("step_delta_"):
            k = key[len("step_delta_"):]
            d = sd[k].detach() - sd0[k]
            # the first Adam step moves every weight by ~lr*sign(g): compare the deltas
            assert (d - golden[key]).abs().max().item() < 3e-5, k


@pytest.mark.parametrize("tag", ["distill_step1_eps", "distill_step2_eps", "distill_step2_v"])
def test_distillation_loss_and_grads_vs_reference_fixture(golden, tag):
    """Progressive distillation (gaussian_diffusion.py:105-154) through the reference's own DiffusionModel.loss with a
    teacher: loss, the four cond_w_embed gradients (simple_unet.py:60-64) and every gradient norm."""
    f = golden[tag]
    sd = {k: v.clone().requires_grad_(True) for k, v in O.random_state_dict(1).items()}
    loss, _, _ = O.training_loss_distill(sd, O.random_state_dict(2), f["x"], f["y"], f["eps"], f["u"], f["cond_w"], f["mode"],
                                         f["T"], f["mean_type"], i=f["i"])
    loss.backward()
    assert abs(loss.item() - f["loss"].item()) <= 1e-6 * abs(f["loss"].item())
    for k, g in f["grads"].items():
        assert ((sd[k].grad - g).abs().max() / g.abs().max()).item() < 1e-5, k
    for k, n in f["gradnorms"].items():
        assert abs(sd[k].grad.norm().item() - n.item()) <= 1e-4 * n.item() + 1e-12, k
    assert all(sd[k].grad is not None for k in sd if k.startswith("cond_w_embed."))


def test_distilled_student_sampling_vs_reference_fixture(golden):
    """With a teacher present the student is conditioned on w = 4*rand(n) and sampled WITHOUT guidance
    (gaussian_diffusion.py:252-255); sampler='teacher_test' runs guided DDIM steps of the teacher (:271-278)."""
    f = golden["distill_step1_eps"]
    sd, sdt = O.random_state_dict(1), O.random_state_dict(2)
    s = f["sample"]
    with torch.no_grad():
        net = lambda z, l, uncond: O.unet_forward(sd, z, l, guide=s["y"], cond_w=s["net_cond_w"])
        zs, xs, es = O.sample_loop(net, s["init"], f["T"], sampler="ddim", mean_type="eps", cond_w=None)
    assert (zs - s["zs"]).abs().max() < 1e-4 and (xs - s["xs"]).abs().max() < 1e-4
    t = f["teacher_test"]
    with torch.no_grad():
        net = lambda z, l, uncond: O.unet_forward(sdt, z, l, guide=(-torch.ones_like(t["y"]) if uncond else t["y"]))
        zs, xs, es = O.sample_loop(net, t["init"], f["T"], sampler="ddim", mean_type="eps", cond_w=t["net_cond_w"])
    assert (zs - t["zs"]).abs().max() < 1e-4 and (xs - t["xs"]).abs().max() < 1e-4


def unpack_keep(entry):
    """fixture["dropout"]["keep"][block] -> bool [B, C, H, W] (bit-packed by oracle/make_golden.py)."""
    import numpy as np
    shape, packed = entry
    n = 1
    for d in shape:
        n *= d
    return torch.from_numpy(np.unpackbits(packed.numpy())[:n].astype(bool)).view(*shape)


def test_dropout_placement_and_scaling_vs_reference_fixture(golden):
    """The reference SimpleUnet(128, dropout=0.3) in train mode, its nn.Dropout keep-decisions recorded by forward hooks:
    the oracle handed the same masks (multiplier keep / (1 - p) after the SiLU of out_layers, simple_unet.py:168-172)
    reproduces the reference's output."""
    f = golden["dropout"]
    sd = O.random_state_dict(1)
    keep = {k: unpack_keep(v) for k, v in f["keep"].items()}
    # eleven blocks drop 30 %; up.seq.3's ResBlock is built without the dropout argument (simple_unet.py:137-139): p = 0
    assert len(keep) == 12 and keep["up.seq.3.0"].all()
    drop = {k: v.float() / (1.0 - f["p"]) for k, v in keep.items() if k != "up.seq.3.0"}
    assert all(abs(v.gt(0).float().mean().item() - (1 - f["p"])) < 0.02 for v in drop.values())
    out = O.unet_forward(sd, f["x"], f["logsnr"], guide=f["y"], drop=drop)
    assert rel(out, f["out"]) < 1e-5
    assert rel(O.unet_forward(sd, f["x"], f["logsnr"], guide=f["y"]), f["out"]) > 1e-2  # without the masks it does not
