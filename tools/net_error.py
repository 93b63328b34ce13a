import sys, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import __graft_entry__ as ge
ge.build()
from generative_models_b200 import lib
from generative_models_b200.model import DiffusionModelB200, default_G
from generative_models_b200.plan import UNetPlan
from oracle import ddpm_oracle as O
golden = torch.load(sys.path[0] + "/tests/golden/reference_fixtures.pt")
m = DiffusionModelB200(default_G(device="cuda:0", mean_type="eps"))
m.net.load_state_dict(O.random_state_dict(1))
p = UNetPlan(m.store, 3, 28, train=False)
p.z_in.copy_(golden["unet_z"].reshape(3, -1)); p.t.copy_(golden["unet_logsnr"]); p.guide.copy_(golden["unet_guide"])
p.forward(m._stream())
o = p.out.cpu().view(3, 1, 28, 28); r = golden["unet_out"]
print("maxrel", ((o - r).abs().max() / r.abs().max()).item(), "rmsrel", ((o - r).pow(2).mean().sqrt() / r.pow(2).mean().sqrt()).item())
for seed in range(4):
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(8, 1, 28, 28, generator=g); l = torch.rand(8, generator=g) * 30 - 15; y = torch.randint(0, 10, (8,), generator=g)
    with torch.no_grad():
        r = O.unet_forward(O.random_state_dict(1), z, l, guide=y)
    p = UNetPlan(m.store, 8, 28, train=False)
    p.z_in.copy_(z.reshape(8, -1)); p.t.copy_(l); p.guide.copy_(y)
    p.forward(m._stream())
    o = p.out.cpu().view(8, 1, 28, 28)
    print(seed, "maxrel", ((o - r).abs().max() / r.abs().max()).item(), "rmsrel", ((o - r).pow(2).mean().sqrt() / r.pow(2).mean().sqrt()).item())
