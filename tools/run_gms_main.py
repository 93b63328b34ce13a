#!/usr/bin/env python
"""Drive the B200 DDPM path with the reference's UNMODIFIED `gms/main.py`.

    python tools/run_gms_main.py --reference-root /path/to/generative_models [--synthetic-mnist] -- \
        --model diffusion_model --bs 512 --epochs 1 --mean_type eps --sampler noisy --timesteps 1000 --eval_heavy 0

What it does (nothing in the reference tree is edited):
  * puts the reference root on sys.path and imports `gms.common`;
  * wraps `common.discover_models` (gms/common.py:38-55) so that the registry key `diffusion_model` resolves to
    `generative_models_b200.model.DiffusionModelB200` (and `diffusion_model_b200` as an alias).  main.py:64,66 looks
    the class up there, reads its `DG` (main.py:67-70), calls `Model(G=G).to(G.device)` (main.py:78) and from then
    on only `loss / train_step / evaluate / sample / save / state_dict / parameters / train / eval`;
  * `--synthetic-mnist`: replaces `common.load_mnist` (gms/common.py:102-132) by a loader of uniform[-1,1] 1x28x28
    images with labels 0..9 (the dataset cannot be downloaded offline);
  * installs a stub `ignite.metrics.FID` when ignite is absent (main.py:10 imports it; only `eval_heavy` uses it).
Everything after `--` is handed to gms.main's own argparse.
"""
import argparse
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def synthetic_mnist(n_train=8192, n_test=1024, seed=0):
    import torch
    from torch.utils import data

    def load_mnist(bs, binarize=True, pad32=False):
        g = torch.Generator().manual_seed(seed)
        side = 32 if pad32 else 28

        def make(n):
            x = torch.rand(n, 1, side, side, generator=g) * 2 - 1
            if binarize:
                x = (x > 0).float()
            y = torch.randint(0, 10, (n,), generator=g)
            return data.TensorDataset(x, y)

        kw = dict(batch_size=bs, shuffle=True, pin_memory=torch.cuda.is_available(), drop_last=True)
        return data.DataLoader(make(n_train), **kw), data.DataLoader(make(n_test), **kw)

    return load_mnist


def install(reference_root, synthetic=False, n_train=8192, n_test=1024):
    """Patch the registry (and optionally the data loader) of the reference's `gms.common`; returns the module."""
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    try:
        import ignite.metrics  # noqa: F401
    except Exception:
        ig = types.ModuleType("ignite")
        igm = types.ModuleType("ignite.metrics")

        class FID:  # only eval_heavy (main.py:108) constructs it
            def __init__(self, *a, **k):
                raise RuntimeError("ignite is not installed: run with --eval_heavy 0")

        igm.FID = FID
        ig.metrics = igm
        sys.modules["ignite"], sys.modules["ignite.metrics"] = ig, igm
    from gms import common

    from generative_models_b200.model import DiffusionModelB200

    stock = common.discover_models

    def discover_models():
        models = stock()
        models["diffusion_model"] = DiffusionModelB200
        models["diffusion_model_b200"] = DiffusionModelB200
        return models

    common.discover_models = discover_models
    if synthetic:
        common.load_mnist = synthetic_mnist(n_train, n_test)
    return common


def default_reference_root():
    """baseline/_ref (the unmodified reference installed by tools/vendor_reference.sh; it travels to the GPU box), else
    $GMS_ROOT, else /root/reference."""
    vend = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isfile(os.path.join(vend, "gms", "main.py")):
        return vend
    return os.environ.get("GMS_ROOT", "/root/reference")


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--reference-root", default=default_reference_root())
    ap.add_argument("--synthetic-mnist", action="store_true")
    ap.add_argument("--n-train", type=int, default=8192, help="size of the synthetic training set")
    ap.add_argument("--n-test", type=int, default=1024)
    ap.add_argument("--git-cwd", action="store_true",
                    help="run from a fresh temp `git init` directory (gms/common.py:74-76 shells out to git rev-parse)")
    if "--" in sys.argv:
        k = sys.argv.index("--")
        own, rest = sys.argv[1:k], sys.argv[k + 1:]
    else:
        own, rest = sys.argv[1:], []
    a = ap.parse_args(own)
    if a.git_cwd:
        import subprocess
        import tempfile
        tmp = tempfile.mkdtemp(prefix="gms_run_")
        env = dict(os.environ, GIT_AUTHOR_NAME="x", GIT_AUTHOR_EMAIL="x@x", GIT_COMMITTER_NAME="x", GIT_COMMITTER_EMAIL="x@x")
        subprocess.check_call(["git", "init", "-q", tmp])
        subprocess.check_call(["git", "-C", tmp, "commit", "-q", "--allow-empty", "-m", "run"], env=env)
        os.chdir(tmp)
    install(a.reference_root, a.synthetic_mnist, a.n_train, a.n_test)
    sys.argv = ["gms.main"] + rest
    import runpy

    runpy.run_module("gms.main", run_name="__main__")


if __name__ == "__main__":
    main()
