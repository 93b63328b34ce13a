"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE ITSELF (CPU, fp32).

Run in the build container only (it imports /root/reference, which does not exist on the GPU box):

    python oracle/make_golden.py

The reference's own tests hold no golden vectors (tests/test_models.py:7-13 only checks an exit code), so these
fixtures — outputs of the unmodified reference modules on seeded inputs — are what pins oracle/ddpm_oracle.py.
Weights are not stored: `ddpm_oracle.random_state_dict(seed)` regenerates them from a CPU generator.
"""
import os
import sys
from functools import partial
from pathlib import Path

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("GMS_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from oracle import ddpm_oracle as O  # noqa: E402


def ref_model(mean_type, sampler, timesteps, sample_cond_w=-1.0, seed=1):
    from gms import common
    from gms.diffusion.diffusion_model import DiffusionModel

    G = common.AttrDict(lr=3e-4, bs=64, device="cpu", pad32=0, weights_from=Path("."), logdir=Path("/tmp"))
    G.update(DiffusionModel.DG)
    G.update(mean_type=mean_type, sampler=sampler, timesteps=timesteps, sample_cond_w=sample_cond_w)
    m = DiffusionModel(G)
    sd = O.random_state_dict(seed)
    m.net.load_state_dict(sd)
    m.eval()
    return m, sd


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    torch.set_num_threads(8)
    fx = {}

    # ---- 1. schedule (diffusion_utils.py:198-201) and the sampler's per-step pair (gaussian_diffusion.py:288-290)
    from gms.diffusion.diffusion_utils import get_logsnr_schedule
    fn = get_logsnr_schedule("cosine", logsnr_min=-20.0, logsnr_max=20.0)
    u = torch.linspace(0, 1, 41)
    fx["sched_u"] = u
    fx["sched_logsnr"] = fn(u)
    T = 8
    fx["sched_T8"] = torch.stack([torch.stack([fn((torch.tensor(i) + 1.0) / T), fn(torch.tensor(i) / T)]) for i in range(T)])

    # ---- 2. UNet forward (simple_unet.py:44-72), guide with a -1 row, all weights randomised
    m, sd = ref_model("eps", "noisy", 4)
    g = torch.Generator().manual_seed(10)
    z = torch.randn(3, 1, 28, 28, generator=g)
    logsnr = torch.tensor([-7.5, 0.3, 11.0])
    guide = torch.tensor([3, -1, 9])
    with torch.no_grad():
        fx["unet_z"], fx["unet_logsnr"], fx["unet_guide"] = z, logsnr, guide
        fx["unet_out"] = m.net(z, logsnr, guide=guide)
        fx["unet_out_noguide"] = m.net(z, logsnr)

    # ---- 3. loss forward + selected gradients (gaussian_diffusion.py:81-100,156-172), the three mean types
    for mt in ("eps", "x", "v"):
        m, sd = ref_model(mt, "noisy", 4)
        g = torch.Generator().manual_seed(20)
        x = torch.rand(4, 1, 28, 28, generator=g) * 2 - 1
        y = torch.tensor([1, -1, 7, 0])
        torch.manual_seed(123)
        m.zero_grad()
        loss, _ = m.loss(x, y)
        loss.backward()
        torch.manual_seed(123)
        eps = torch.randn(x.shape)
        uu = torch.rand(4)
        fx[f"loss_{mt}_x"], fx[f"loss_{mt}_y"], fx[f"loss_{mt}_eps"], fx[f"loss_{mt}_u"] = x, y, eps, uu
        fx[f"loss_{mt}"] = loss.detach()
        grads = {k: p.grad for k, p in m.net.named_parameters() if p.grad is not None}
        fx[f"loss_{mt}_gradnorms"] = {k: v.norm() for k, v in grads.items()}
        for k in ("out.2.weight", "down.seq.0.conv.weight", "turn.in_layers.0.weight", "time_embed.0.bias",
                  "up.seq.3.0.skip_connection.bias", "guide_embed.2.bias", "up.seq.5.emb_layers.1.bias"):
            fx[f"loss_{mt}_grad_{k}"] = grads[k].clone()

    # ---- 4. sampler trajectories (gaussian_diffusion.py:245-296): noisy+guided, ddim+unguided, v/eps
    for tag, mt, sampler, guided, T in (("noisy_guided", "eps", "noisy", True, 4), ("ddim_plain", "v", "ddim", False, 3),
                                        ("noisy_plain", "eps", "noisy", False, 3)):
        m, sd = ref_model(mt, sampler, T)
        g = torch.Generator().manual_seed(30)
        init = torch.randn(2, 1, 28, 28, generator=g)
        y = torch.tensor([4, 8])
        torch.manual_seed(77)
        with torch.no_grad():
            zs, xs, es = m.diffusion.sample(net=partial(m.net, guide=y), init_x=init, cond_w=0.5 if guided else None)
        torch.manual_seed(77)
        cw = 4.0 * torch.rand(2) if guided else None
        noises = [torch.randn(init.shape) for _ in range(T)] if sampler == "noisy" else None
        fx[f"samp_{tag}"] = dict(init=init, y=y, cond_w=cw, noises=noises, zs=zs, xs=xs, es=es, mean_type=mt,
                                 sampler=sampler, T=T)

    # ---- 5. one Adam step through DiffusionModel.train_step (diffusion_model.py:63-74) on CPU (autocast inert)
    m, sd = ref_model("eps", "noisy", 4)
    m.train()
    g = torch.Generator().manual_seed(40)
    x = torch.rand(4, 1, 28, 28, generator=g) * 2 - 1
    y = torch.tensor([5, 2, 2, 9])
    torch.manual_seed(5)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        metrics = m.train_step(x, y.clone())
    torch.manual_seed(5)
    drop = torch.rand(4) < 0.1
    eps = torch.randn(x.shape)
    uu = torch.rand(4)
    y2 = y.clone()
    y2[drop] = -1
    fx["step_x"], fx["step_y"], fx["step_eps"], fx["step_u"] = x, y2, eps, uu
    fx["step_loss"] = metrics["loss"].detach()
    new = dict(m.net.named_parameters())
    for k in ("out.2.bias", "down.seq.1.in_layers.2.weight", "time_embed.2.weight", "up.seq.6.out_layers.3.weight"):
        fx[f"step_delta_{k}"] = (new[k].detach() - sd[k]).clone()

    # ---- 6. the reference's own initialisation at seed 0 (simple_unet.py:16-42: RNG consumption order)
    from gms.diffusion.simple_unet import SimpleUnet
    torch.manual_seed(0)
    net0 = SimpleUnet(128, 0.0)
    sd0 = net0.state_dict()
    fx["init_seed0_checks"] = {k: sd0[k].permute(0, 2, 3, 1).contiguous().flatten()[:64].clone() if sd0[k].ndim == 4
                               else sd0[k].flatten()[:64].clone()
                               for k in ("time_embed.0.weight", "guide_embed.2.bias", "down.seq.0.conv.weight",
                                         "down.seq.4.in_layers.2.weight", "turn.emb_layers.1.weight",
                                         "up.seq.3.0.skip_connection.weight", "up.seq.6.in_layers.2.bias", "out.2.weight")}

    # ---- 7. progressive distillation (gaussian_diffusion.py:105-154, diffusion_model.py:34-43): step1 and step2 losses,
    # all four cond_w_embed gradients and a few more, through the reference's own DiffusionModel.loss with a teacher
    import tempfile
    from gms import common
    from gms.diffusion.diffusion_model import DiffusionModel
    tdir = tempfile.mkdtemp()
    teacher_sd = {"net." + k: v for k, v in O.random_state_dict(2).items()}
    torch.save(teacher_sd, os.path.join(tdir, "teacher.pt"))
    for mode, mt, T in (("step1", "eps", 8), ("step2", "eps", 4), ("step2", "v", 4)):
        G = common.AttrDict(lr=3e-4, bs=64, device="cpu", pad32=0, weights_from=Path("."), logdir=Path("/tmp"))
        G.update(DiffusionModel.DG)
        G.update(mean_type=mt, sampler="ddim", timesteps=T, teacher_path=Path(os.path.join(tdir, "teacher.pt")), teacher_mode=mode)
        m = DiffusionModel(G)
        assert m.teacher_net is not None
        m.net.load_state_dict(O.random_state_dict(1))  # student != teacher, so the targets are not trivially met
        m.eval()
        g = torch.Generator().manual_seed(50)
        x = torch.rand(4, 1, 28, 28, generator=g) * 2 - 1
        y = torch.tensor([3, -1, 0, 8])
        torch.manual_seed(321)
        m.zero_grad()
        loss, _ = m.loss(x, y)
        loss.backward()
        torch.manual_seed(321)
        eps = torch.randn(x.shape)
        if mode == "step2":
            ii = torch.randint(T, (4,))
            uu = (ii + 1).to(x.dtype) / T
        else:
            ii = None
            uu = torch.rand(4)
        cw = 4.0 * torch.rand_like(uu)
        tag = f"distill_{mode}_{mt}"
        fx[tag] = dict(x=x, y=y, eps=eps, u=uu, i=ii, cond_w=cw, T=T, mode=mode, mean_type=mt, loss=loss.detach())
        grads = {k: p.grad for k, p in m.net.named_parameters() if p.grad is not None}
        fx[tag]["gradnorms"] = {k: v.norm() for k, v in grads.items()}
        fx[tag]["grads"] = {k: grads[k].clone() for k in ("cond_w_embed.0.weight", "cond_w_embed.0.bias", "cond_w_embed.2.weight",
                                                         "cond_w_embed.2.bias", "out.2.weight", "time_embed.0.bias",
                                                         "turn.in_layers.0.weight")}
        assert all(p.grad is None for p in m.teacher_net.parameters())
        # student sampling with a teacher present (gaussian_diffusion.py:252-255): net conditioned on w, no guidance
        if mode == "step1":
            init = torch.randn(2, 1, 28, 28, generator=g)
            ys = torch.tensor([4, 8])
            torch.manual_seed(78)
            with torch.no_grad():
                zs, xs, es = m.diffusion.sample(net=partial(m.net, guide=ys), init_x=init, cond_w=0.5)
            torch.manual_seed(78)
            fx[tag]["sample"] = dict(init=init, y=ys, net_cond_w=4.0 * torch.rand(2), zs=zs, xs=xs, es=es)
            # sampler='teacher_test' (:271-278): guided DDIM steps of the teacher itself
            m.diffusion.sampler = "teacher_test"
            torch.manual_seed(79)
            with torch.no_grad():
                zs, xs, es = m.diffusion.sample(net=partial(m.net, guide=ys), init_x=init, cond_w=0.5)
            torch.manual_seed(79)
            fx[tag]["teacher_test"] = dict(init=init, y=ys, net_cond_w=4.0 * torch.rand(2), zs=zs, xs=xs, es=es)

    # ---- 8. dropout (simple_unet.py:168-172: GroupNorm -> SiLU -> nn.Dropout(p) -> conv in ResBlock.out_layers).  The
    # reference SimpleUnet with p = 0.3 in train mode; forward hooks on its nn.Dropout modules record which elements were
    # kept, so the oracle (and the device tests) can be handed the same masks.  Stored bit-packed.
    import numpy as np
    from torch import nn
    from gms.diffusion.simple_unet import SimpleUnet
    p_drop = 0.3
    net = SimpleUnet(128, p_drop)
    net.load_state_dict(O.random_state_dict(1))
    net.train()
    g = torch.Generator().manual_seed(2024)
    xd = torch.randn(2, 1, 28, 28, generator=g)
    ld = torch.tensor([1.5, -3.0])
    yd = torch.tensor([7, -1])
    keep = {}

    def hook(name):
        def fn(mod, inp, out):
            keep[name] = (out != 0) | (inp[0] == 0)  # an input that is exactly 0 says nothing: count it as kept
        return fn

    handles = [mod.register_forward_hook(hook(name[:-len(".out_layers.2")]))
               for name, mod in net.named_modules() if isinstance(mod, nn.Dropout)]
    torch.manual_seed(4242)
    with torch.no_grad():
        out_d = net(xd, ld, guide=yd)
    for h in handles:
        h.remove()
    assert len(keep) == 12, sorted(keep)
    fx["dropout"] = dict(p=p_drop, x=xd, logsnr=ld, y=yd, out=out_d,
                         keep={k: (tuple(v.shape), torch.from_numpy(np.packbits(v.numpy().reshape(-1)))) for k, v in keep.items()})

    torch.save(fx, os.path.join(out_dir, "reference_fixtures.pt"))
    print("wrote", os.path.join(out_dir, "reference_fixtures.pt"),
          os.path.getsize(os.path.join(out_dir, "reference_fixtures.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
