"""Locate and import the UNMODIFIED reference (matwilso/generative_models, package `gms`) — TEST INFRASTRUCTURE ONLY.

Only tests/, bench.py's cpu_baseline / `--impl reference` legs and tools/gpu_torch_baseline.py use this module; the
product package never does.

Where the reference comes from: `/root/reference` exists only in the build container.  `tools/vendor_reference.sh`
(also a step of `__graft_entry__.build()`) pip-installs it, unmodified, into the git-ignored `baseline/_ref/`
(`pip install --no-index --no-deps --target baseline/_ref <copy of /root/reference>`), which travels to the GPU box
with the gpurun snapshot.  Nothing under baseline/_ref is committed.

Shims needed to import `gms.main` offline (none of them touches a reference file):
  * `ignite.metrics.FID` (gms/main.py:10) is imported but only used by eval_heavy → a stub module when ignite is absent;
  * `common.load_mnist` (gms/common.py:102-132) downloads MNIST → replaced by a synthetic loader of the same shapes;
  * `common.dump_logger` (gms/common.py:74-76) shells out to git → callers run from a temp `git init` directory.
"""
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = [os.path.join(ROOT, "baseline", "_ref"), os.environ.get("GMS_REFERENCE", "/root/reference")]


def find_root():
    """Directory that contains the reference's `gms` package, or None."""
    for c in CANDIDATES:
        if c and os.path.isfile(os.path.join(c, "gms", "diffusion", "diffusion_model.py")):
            return c
    return None


def available():
    return find_root() is not None


def install(synthetic_mnist=True, n_train=8192, n_test=1024):
    """Put the reference on sys.path, install the offline shims, return its `gms.common` module."""
    root = find_root()
    if root is None:
        raise RuntimeError("reference not found: run tools/vendor_reference.sh in the build container "
                           "(installs /root/reference into baseline/_ref)")
    if root not in sys.path:
        sys.path.insert(0, root)
    try:
        import ignite.metrics  # noqa: F401
    except Exception:
        ig, igm = types.ModuleType("ignite"), types.ModuleType("ignite.metrics")

        class FID:  # only eval_heavy (gms/main.py:108) constructs it
            def __init__(self, *a, **k):
                raise RuntimeError("ignite is not installed: run with --eval_heavy 0")

        igm.FID = FID
        ig.metrics = igm
        sys.modules["ignite"], sys.modules["ignite.metrics"] = ig, igm
    from gms import common

    if synthetic_mnist:
        common.load_mnist = synthetic_loader(n_train, n_test)
    return common


def synthetic_loader(n_train=8192, n_test=1024, seed=0):
    """Stand-in for gms/common.py:102-132 load_mnist: uniform[-1,1] 1x28x28 images (the `binarize=0` preprocessing is
    2x-1, common.py:108-109), labels 0..9, same DataLoader settings minus the worker processes."""
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

        kw = dict(batch_size=bs, shuffle=True, drop_last=True)
        return data.DataLoader(make(n_train), **kw), data.DataLoader(make(n_test), **kw)

    return load_mnist


def reference_model(device, **over):
    """The reference's DiffusionModel built the way gms/main.py:43-78 builds it (main's DG, then the model's DG, then
    overrides), on `device`."""
    from pathlib import Path

    install()
    from gms import common
    from gms.diffusion.diffusion_model import DiffusionModel

    G = common.AttrDict(lr=3e-4, bs=64, device=device, pad32=0, weights_from=Path("."), logdir=Path("/tmp"),
                        model="diffusion_model", epochs=1, save_n=5, mode="train", skip_training=0)
    G.update(DiffusionModel.DG)
    G.update(over)
    return DiffusionModel(G=G).to(device), G
