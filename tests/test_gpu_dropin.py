"""-m gpu: the drop-in boundary on hardware and the host-side step machinery.

  * the reference's UNMODIFIED gms/main.py (from baseline/_ref, installed by tools/vendor_reference.sh) drives
    DiffusionModelB200 through tools/run_gms_main.py: one epoch of 2 batches, T = 8, evaluate() + save();
  * the whole train step is ONE captured graph that is not re-captured from step to step (the c_float lr
    comparison used to re-capture the Adam graph every step), and follows an lr change through the device-resident
    {lr, grad_scale};
  * device-side label drop (G.device_label_drop=1): drop rate, and that -1 rows take the unconditional embedding.
"""
import os
import subprocess
import sys

import pytest
import torch

from oracle import ddpm_oracle as O
from oracle import reference_runtime as R

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make_model(**over):
    from generative_models_b200.model import DiffusionModelB200, default_G
    m = DiffusionModelB200(default_G(device=DEV, **over))
    m.net.load_state_dict(O.random_state_dict(1))
    return m


@pytest.mark.skipif(not R.available(), reason="reference not vendored (run tools/vendor_reference.sh in the build container)")
def test_unmodified_gms_main_drives_the_b200_model(built_lib, tmp_path):
    logdir = tmp_path / "logs"
    cmd = [sys.executable, os.path.join(ROOT, "tools", "run_gms_main.py"), "--reference-root", R.find_root(),
           "--synthetic-mnist", "--n-train", "128", "--n-test", "64", "--git-cwd", "--",
           "--model", "diffusion_model", "--bs", "64", "--epochs", "1", "--mean_type", "eps", "--sampler", "noisy",
           "--timesteps", "8", "--eval_heavy", "0", "--logdir", str(logdir), "--device", "cuda"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    out = r.stdout
    assert "num_vars 6033665" in out
    assert "diffusion_model/train/loss" in out and "diffusion_model/test/loss" in out and "SAVED MODEL" in out
    # an explicit --logdir is used as given (main.py:70 only defaults it to logdir/<model>)
    sd = torch.load(logdir / "model.pt", map_location="cpu")
    assert len(sd) == 160 and all(k.startswith("net.") for k in sd)
    assert (logdir / "hps.yaml").exists()
    losses = [float(ln.split()[-1]) for ln in out.splitlines() if ln.startswith("diffusion_model/train/loss ")]
    assert losses and all(0 < v < 10 for v in losses), losses


def test_train_step_is_one_graph_captured_once(built_lib):
    m = make_model(mean_type="eps", lr=3e-4)
    g = torch.Generator().manual_seed(0)
    B = 8
    x = (torch.rand(B, 1, 28, 28, generator=g) * 2 - 1).to(DEV)
    y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    for _ in range(2):
        m.train_step(x, y.clone())
    b = m._plan(B, "train")
    keys = [k for k in b.graphs if isinstance(k, tuple) and k[0] == "step"]
    assert len(keys) == 1, list(b.graphs)
    g0 = b.graphs[keys[0]]
    before = m.store.flat.clone()
    for _ in range(3):
        m.train_step(x, y.clone())
    assert b.graphs[keys[0]] is g0 and len([k for k in b.graphs if isinstance(k, tuple) and k[0] == "step"]) == 1
    d1 = (m.store.flat - before).abs().max().item()
    assert 0 < d1 < 4 * 3e-4
    # an lr change reaches the captured Adam through device memory: no re-capture, bigger step
    m.G.lr = 3e-2
    before = m.store.flat.clone()
    m.train_step(x, y.clone())
    assert b.graphs[keys[0]] is g0
    d2 = (m.store.flat - before).abs().max().item()
    assert d2 > 3e-3, (d1, d2)
    assert int(m.adam_step.item()) == 6


def test_graphed_step_draws_like_eager(built_lib):
    """eps / u drawn inside the captured step continue torch's Philox stream exactly like eager draws."""
    outs = []
    for graph in (0, 1):
        m = make_model(mean_type="eps", graph=graph, cf_drop_prob=0.0)
        B = 4
        x = torch.zeros(B, 1, 28, 28, device=DEV)
        y = torch.zeros(B, dtype=torch.long, device=DEV)
        torch.manual_seed(99)
        for _ in range(3):  # eager warm-up, capture, replay
            m.train_step(x, y.clone())
        b = m._plan(B, "train")
        outs.append((b.eps.clone(), b.u.clone()))
    torch.manual_seed(99)
    for _ in range(3):
        eps = torch.randn(B, 1, 28, 28, device=DEV)
        u = torch.rand(B, device=DEV)
    for e, uu in outs:
        assert torch.equal(e.view_as(eps), eps) and torch.equal(uu, u)


def test_device_label_drop(built_lib):
    m = make_model(mean_type="eps", device_label_drop=1, cf_drop_prob=0.25)
    B = 4096 // 8
    g = torch.Generator().manual_seed(0)
    x = (torch.rand(B, 1, 28, 28, generator=g) * 2 - 1).to(DEV)
    y = torch.randint(0, 10, (B,), generator=g).to(DEV)
    y_in = y.clone()
    dropped = 0
    for _ in range(4):
        m.train_step(x, y_in)
        guide = m._plan(B, "train").plan.guide
        assert torch.equal(guide[guide != -1], y[guide != -1])
        dropped += int((guide == -1).sum())
    assert torch.equal(y_in, y)  # device mode leaves the caller's labels alone
    rate = dropped / (4 * B)
    assert 0.2 < rate < 0.3, rate
    # a -1 row takes the unconditional embedding: same output as guide=None rows of the oracle
    m2 = make_model(mean_type="eps")
    sd = O.random_state_dict(1)
    z = torch.randn(3, 1, 28, 28, generator=g)
    l = torch.tensor([-3.0, 0.5, 4.0])
    p = m2._plan(3, "loss").plan
    p.z_in.copy_(z.reshape(3, -1).to(DEV))
    p.t.copy_(l.to(DEV))
    p.guide.copy_(torch.tensor([-1, 4, -1], device=DEV))
    p.forward(m2._stream())
    with torch.no_grad():
        ref = O.unet_forward(sd, z, l, guide=torch.tensor([-1, 4, -1]))
        ref_none = O.unet_forward(sd, z, l, guide=None)
    ours = p.out.view(3, 1, 28, 28).cpu()
    rms = lambda a, b: ((a - b).pow(2).mean().sqrt() / b.pow(2).mean().sqrt()).item()
    assert rms(ours, ref) < 1e-3
    assert rms(ours[0], ref_none[0]) < 1e-3 and rms(ours[2], ref_none[2]) < 1e-3


def test_evaluate_follows_sample_cond_w(built_lib):
    """diffusion_model.py:99-103 passes cond_w=None, gaussian_diffusion.py:257 then applies G.sample_cond_w when it is
    not -1: evaluate()'s 25 samples are guided (one 2n-row forward per step) exactly in that case."""
    class W:
        def __init__(self):
            self.img = {}

        def add_image(self, tag, img, epoch):
            self.img[tag] = img.clone()

        def add_video(self, *a, **k):
            pass

    grids = []
    for scw in (-1.0, 0.5):
        m = make_model(mean_type="eps", sampler="ddim", timesteps=4, sample_cond_w=scw)
        w = W()
        m.evaluate(w, None, None, 0)
        grids.append(w.img["samples"])
        assert ("sample_g" in [k[0] for k in m._plans]) == (scw != -1.0)
    assert grids[0].shape == (1, 140, 140) and not torch.equal(grids[0], grids[1])
