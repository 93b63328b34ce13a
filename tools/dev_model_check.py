"""Developer probe: accuracy of the whole path vs the CPU oracle / reference fixtures, printed (no asserts)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ddpm_oracle as O  # noqa: E402
from generative_models_b200.model import DiffusionModelB200, default_G  # noqa: E402
from generative_models_b200.plan import UNetPlan  # noqa: E402

DEV = "cuda:0"


def rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()


def rms(a, b):
    return ((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt().clamp_min(1e-30)).item()


def main():
    golden = torch.load(os.path.join(ROOT, "tests", "golden", "reference_fixtures.pt"), weights_only=False)
    m = DiffusionModelB200(default_G(device=DEV, mean_type="eps"))
    sd0 = O.random_state_dict(1)
    m.net.load_state_dict(sd0)
    p = UNetPlan(m.store, 3, 28, train=True)
    p.z_in.copy_(golden["unet_z"].reshape(3, -1))
    p.t.copy_(golden["unet_logsnr"])
    p.guide.copy_(golden["unet_guide"])
    p.forward(m._stream())
    out = p.out.cpu().view(3, 1, 28, 28)
    print("unet fwd vs reference fixture: max-rel %.3e rms-rel %.3e" % (rel(out, golden["unet_out"]), rms(out, golden["unet_out"])))
    taps = {}
    with torch.no_grad():
        ref = O.unet_forward(sd0, golden["unet_z"], golden["unet_logsnr"], guide=golden["unet_guide"], taps=taps)
    print("emb: %.3e" % rel(p.emb.cpu(), taps["emb"]))
    sv = p.saved
    for nm, t, r in (("down(d6)", sv["d6"], taps["down"]), ("turn", sv["t"], taps["turn"]), ("up(u6)", sv["u6"], taps["up"])):
        B, HW, Cc = t.shape
        H = int(HW ** 0.5)
        tt = t.cpu().view(B, H, H, Cc).permute(0, 3, 1, 2)
        print(f"{nm}: max-rel {rel(tt, r):.3e} rms-rel {rms(tt, r):.3e}")
    # gradients
    g = torch.Generator().manual_seed(77)
    B = 6
    x = torch.rand(B, 1, 28, 28, generator=g) * 2 - 1
    y = torch.tensor([0, 5, -1, 9, 3, -1])
    eps = torch.randn(B, 1, 28, 28, generator=g)
    u = torch.rand(B, generator=g) * 0.8 + 0.1
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    loss, per, aux = O.training_loss(sd, x, y, eps, u, "eps")
    loss.backward()
    m.store.grad.zero_()
    l2, b = m.loss_and_grads(x.to(DEV), y.to(DEV), eps=eps.to(DEV), u=u.to(DEV))
    print("loss ours %.6f oracle %.6f" % (l2.item(), loss.item()))
    print("net out: max-rel %.3e rms-rel %.3e" % (rel(b.plan.out.cpu().view(B, 1, 28, 28), aux["out"].detach()),
                                                   rms(b.plan.out.cpu().view(B, 1, 28, 28), aux["out"].detach())))
    print("per-sample loss rel %.3e" % rel(b.loss_b.cpu(), per.detach()))
    errs = []
    for k, prm in m.net.named_parameters():
        r = sd[k].grad
        if r is None:
            continue
        errs.append((rel(prm.grad.cpu(), r), rms(prm.grad.cpu(), r), k))
    errs.sort(reverse=True)
    print("worst grads (max-rel, rms-rel):")
    for e in errs[:12]:
        print("   %.3e %.3e %s" % e)
    print("median grad max-rel %.3e" % errs[len(errs) // 2][0])
    # ---- backward alone: linear loss sum(out * r) so the (ill-conditioned, clipped) diffusion loss is out of the way
    r = torch.randn(B, 1, 28, 28, generator=g)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    logsnr = O.logsnr_cosine(u)
    z = O.q_sample(x, logsnr, eps)
    (O.unet_forward(sd, z, logsnr, guide=y) * r).sum().backward()
    m.store.grad.zero_()
    pl = b.plan
    st = m._stream()
    pl.forward(st)
    pl.dout.copy_(r.reshape(B, -1))
    pl.backward(st)
    errs = []
    for k, prm in m.net.named_parameters():
        rr = sd[k].grad
        if rr is None:
            continue
        errs.append((rel(prm.grad.cpu(), rr), rms(prm.grad.cpu(), rr), k))
    errs.sort(reverse=True)
    print("linear-loss backward, worst grads (max-rel, rms-rel):")
    for e in errs[:12]:
        print("   %.3e %.3e %s" % e)
    print("median %.3e" % errs[len(errs) // 2][0])


if __name__ == "__main__":
    main()
