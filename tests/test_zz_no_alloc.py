"""-m gpu: `b200_gn_bwd_static` — the GroupNorm(+SiLU) backward that allocates nothing — against `b200_gn_bwd` with a
caller-allocated workspace, through the C ABI, on two calls in a row.

The file is named so that pytest collects it LAST.  It makes no assumption about what ran
before it in the process: it passes alone in a fresh process and after the rest of the suite.

HW = 4096 (64 x 64) takes the streaming path, the one that reduces through the N*64-float scratch; HW = 784 takes the
shared-memory-resident kernel (scratch unused) and must behave the same.  Tolerance: the 1e-4 of max|ref| the GroupNorm
backward tests use (tests/test_gpu_convs.py::test_gn_silu_fwd_bwd) — both entry points run the same kernels, the
difference is the order of the floating-point atomics.
"""
import pytest
import torch

from tests.gpu_util import ptr, rel, stream

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-4


@pytest.fixture(scope="module")
def lib(built_lib):
    built_lib.load().b200_init(0)
    return built_lib


def call_static(lib, b):
    L = lib.load()
    rc = L.b200_gn_bwd_static(lib.C.byref(b), stream())
    if rc != 0:
        raise lib.B200Error(f"b200_gn_bwd_static failed (rc={rc}): {L.b200_last_error().decode()}")


def make_case(lib, N, HW, C2, seed):
    g = torch.Generator().manual_seed(seed)
    Cc = 128 + C2
    t = {
        "x1": (torch.randn(N, HW, 128, generator=g) * 2 + 0.5).to(DEV),
        "x2": (torch.randn(N, HW, C2, generator=g) * 2 + 0.5).to(DEV) if C2 else None,
        "gamma": (1 + 0.2 * torch.randn(Cc, generator=g)).to(DEV),
        "beta": (0.1 * torch.randn(Cc, generator=g)).to(DEV),
        "dy": torch.randn(N, HW, Cc, generator=g).to(DEV),
        "add1": torch.randn(N, HW, Cc, generator=g).to(DEV),
        "stats": torch.zeros(N, 64, device=DEV),
        "y": torch.empty(N, HW, Cc, device=DEV),
    }
    a = lib.GnArgs(x=ptr(t["x1"]), x2=ptr(t["x2"]), stats=ptr(t["stats"]), gamma=ptr(t["gamma"]), beta=ptr(t["beta"]),
                   y=ptr(t["y"]), N=N, HW=HW, C1=128, C2=C2, x_ld=0, x2_ld=0, eps=1e-5, silu=1, round_tf32=0)
    lib.call("b200_gn_fwd", a, stream())
    return t, a, Cc


def outputs(N, HW, Cc):
    return {"dx": torch.full((N, HW, Cc), float("nan"), device=DEV), "dgamma": torch.zeros(Cc, device=DEV),
            "dbeta": torch.zeros(Cc, device=DEV)}


def bwd_args(lib, a, t, o, Cc):
    return lib.GnBwdArgs(fwd=a, dy=ptr(t["dy"]), dx=ptr(o["dx"]), add1=ptr(t["add1"]), add2=None, add1_ld=Cc, add2_ld=0,
                         add_c=Cc, dgamma=ptr(o["dgamma"]), dbeta=ptr(o["dbeta"]))


@pytest.mark.parametrize("N,HW,C2", [(3, 4096, 0), (2, 4096, 128), (3, 784, 0)])
def test_static_scratch_gn_bwd_matches_workspace_entry_point(lib, N, HW, C2):
    # two different problems back to back: the second call finds the first one's sums in the shared scratch
    for k, n in enumerate((N, max(1, N - 1))):
        t, a, Cc = make_case(lib, n, HW, C2, seed=1000 * k + HW + C2)
        want, got = outputs(n, HW, Cc), outputs(n, HW, Cc)
        bw = bwd_args(lib, a, t, want, Cc)
        ws = torch.zeros(lib.load().b200_gn_bwd_workspace(lib.C.byref(bw)) // 4, device=DEV)
        lib.call("b200_gn_bwd", bw, stream(), ws=ptr(ws), ws_bytes=ws.numel() * 4)
        bg = bwd_args(lib, a, t, got, Cc)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        call_static(lib, bg)
        torch.cuda.synchronize()
        if k == 1:  # (the first call of a process may load the kernels' module; the second must leave memory untouched)
            assert torch.cuda.mem_get_info()[0] == free_before
        for key in want:
            assert torch.isfinite(got[key]).all(), key
            assert rel(got[key].cpu(), want[key].cpu()) < TOL, (k, key)


def test_static_scratch_gn_bwd_rejects_more_samples_than_the_scratch_holds(lib):
    t, a, Cc = make_case(lib, 1, 4096, 0, seed=7)
    o = outputs(1, 4096, Cc)
    b = bwd_args(lib, a, t, o, Cc)
    b.fwd.N = 4096 + 1  # B200_GN_BWD_STATIC_MAX_N + 1: rejected before anything is launched
    with pytest.raises(lib.B200Error, match="exceeds the static scratch"):
        call_static(lib, b)
    torch.cuda.synchronize()
    assert torch.isnan(o["dx"]).all()
