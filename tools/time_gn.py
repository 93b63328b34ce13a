"""Time b200_gn_fwd / b200_gn_bwd alone at the train-step shapes and print achieved HBM GB/s (algorithmic bytes).

    python tools/time_gn.py [--batch 512]          # knobs: B200_GN_CS, B200_GN_THREADS, B200_GN_STREAM
"""
import argparse
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.build()
from generative_models_b200 import lib  # noqa: E402


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def bf16_cases(N, dev, st):
    """The GroupNorm backward launches of the bf16 step (plan.py _resblock_bwd): `out_layers.0` — x, dy bf16, dx bf16
    only, 128 channels — and `in_layers.0` — x fp32, dy bf16, dx fp32 (+ one add tensor) and a bf16 copy of its first
    128 channels.  Algorithmic bytes per element: 6 and 4 + 2 + 4 + 4 (+ 2 on 128 of the channels)."""
    tag = ""
    for HW in (784, 196):
        for kind, Cc in (("out_layers", 128), ("in_layers", 128), ("in_layers", 256)):
            per = N * HW * Cc
            nset = max(2, int(600e6 // (per * 12)) + 1)
            sets = []
            gamma, beta = torch.randn(Cc, device=dev), torch.randn(Cc, device=dev)
            dg, db = torch.zeros(Cc, device=dev), torch.zeros(Cc, device=dev)
            for _ in range(nset):
                stats = torch.zeros(N, 64, device=dev)
                stats.view(N, 32, 2)[..., 1] = 1.0
                dy = torch.randn(N, HW, Cc, device=dev).bfloat16()
                dxb = torch.empty(N, HW, 128, device=dev, dtype=torch.bfloat16)
                if kind == "out_layers":
                    x = torch.randn(N, HW, Cc, device=dev).bfloat16()
                    f = lib.GnArgs(x=ptr(x), x2=None, stats=ptr(stats), gamma=ptr(gamma), beta=ptr(beta), y=None, N=N, HW=HW,
                                   C1=Cc, C2=0, x_ld=0, x2_ld=0, eps=1e-5, silu=1, round_tf32=1, y_dtype=1, x_dtype=1)
                    b = lib.GnBwdArgs(fwd=f, dy=ptr(dy), dx=None, add1=None, add2=None, add1_ld=0, add2_ld=0, add_c=Cc,
                                      dgamma=ptr(dg), dbeta=ptr(db), round_dx=1, dx_bf16=ptr(dxb), dx_bf16_c=128, dy_dtype=1)
                    keep = (x, dy, dxb, stats)
                else:
                    x1 = torch.randn(N, HW, 128, device=dev)
                    x2 = torch.randn(N, HW, 128, device=dev) if Cc == 256 else None
                    dx = torch.empty(N, HW, Cc, device=dev)
                    add1 = torch.randn(N, HW, Cc, device=dev)
                    f = lib.GnArgs(x=ptr(x1), x2=ptr(x2), stats=ptr(stats), gamma=ptr(gamma), beta=ptr(beta), y=None, N=N,
                                   HW=HW, C1=128, C2=Cc - 128, x_ld=0, x2_ld=0, eps=1e-5, silu=1, round_tf32=1, y_dtype=1)
                    b = lib.GnBwdArgs(fwd=f, dy=ptr(dy), dx=ptr(dx), add1=ptr(add1), add2=None, add1_ld=Cc, add2_ld=0,
                                      add_c=Cc, dgamma=ptr(dg), dbeta=ptr(db), round_dx=1, dx_bf16=ptr(dxb), dx_bf16_c=128,
                                      dy_dtype=1)
                    keep = (x1, x2, dy, dx, add1, dxb, stats)
                sets.append((b, keep))
            k = [0]

            def fn():
                k[0] = (k[0] + 1) % nset
                lib.call("b200_gn_bwd", sets[k[0]][0], st)
            us = timeit(fn)
            nbytes = per * 6 if kind == "out_layers" else per * 14 + N * HW * 128 * 2
            print(f"{tag} gn_bwd bf16 {kind:10s} HW{HW} C{Cc}: {us:8.1f} us  {nbytes / us / 1e3:7.0f} GB/s")
            del sets
            torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--bf16", type=int, default=0, help="1: time the bf16-step backward launches instead")
    a = ap.parse_args()
    if a.bf16:
        lib.load().b200_init(0)
        return bf16_cases(a.batch, "cuda:0", C.c_void_p(torch.cuda.current_stream().cuda_stream))
    dev = "cuda:0"
    lib.load().b200_init(0)
    N = a.batch
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    tag = " ".join(f"{k}={os.environ[k]}" for k in ("B200_GN_CS", "B200_GN_THREADS", "B200_GN_STREAM") if k in os.environ)
    for HW in (784, 196, 49):
        for C2 in (0, 128):
            Cc = 128 + C2
            # several rotating buffer sets so that nothing is served from the 126 MB L2
            nset = max(2, int(600e6 // (N * HW * Cc * 4 * 3)) + 1)
            sets = []
            for _ in range(nset):
                x1 = torch.randn(N, HW, 128, device=dev)
                x2 = torch.randn(N, HW, 128, device=dev) if C2 else None
                y = torch.empty(N, HW, Cc, device=dev)
                dy = torch.randn(N, HW, Cc, device=dev)
                dx = torch.empty(N, HW, Cc, device=dev)
                add1 = torch.randn(N, HW, Cc, device=dev)
                stats = torch.zeros(N, 64, device=dev)
                sets.append((x1, x2, y, dy, dx, add1, stats))
            gamma, beta = torch.randn(Cc, device=dev), torch.randn(Cc, device=dev)
            dg, db = torch.zeros(Cc, device=dev), torch.zeros(Cc, device=dev)
            ws = torch.zeros(N * 64, device=dev)
            fa, ba, ba1 = [], [], []
            for x1, x2, y, dy, dx, add1, stats in sets:
                f = lib.GnArgs(x=ptr(x1), x2=ptr(x2), stats=ptr(stats), gamma=ptr(gamma), beta=ptr(beta), y=ptr(y), N=N,
                               HW=HW, C1=128, C2=C2, x_ld=0, x2_ld=0, eps=1e-5, silu=int(os.environ.get("GN_SILU", 1)), round_tf32=1)
                fa.append(f)
                ba.append(lib.GnBwdArgs(fwd=f, dy=ptr(dy), dx=ptr(dx), add1=None, add2=None, add1_ld=0, add2_ld=0,
                                        add_c=Cc, dgamma=ptr(dg), dbeta=ptr(db), round_dx=1))
                ba1.append(lib.GnBwdArgs(fwd=f, dy=ptr(dy), dx=ptr(dx), add1=ptr(add1), add2=None, add1_ld=Cc, add2_ld=0,
                                         add_c=Cc, dgamma=ptr(dg), dbeta=ptr(db), round_dx=1))
            k = [0]

            def run(name, args, use_ws=False):
                def fn():
                    k[0] = (k[0] + 1) % nset
                    lib.call(name, args[k[0]], st, ws=ptr(ws) if use_ws else None, ws_bytes=ws.numel() * 4 if use_ws else 0)
                return timeit(fn)

            el = N * HW * Cc * 4
            us = run("b200_gn_fwd", fa)
            print(f"{tag} gn_fwd  HW{HW} C{Cc}: {us:8.1f} us  {2 * el / us / 1e3:7.0f} GB/s")
            us = run("b200_gn_bwd", ba, True)
            print(f"{tag} gn_bwd  HW{HW} C{Cc}: {us:8.1f} us  {3 * el / us / 1e3:7.0f} GB/s")
            us = run("b200_gn_bwd", ba1, True)
            print(f"{tag} gn_bwd+ HW{HW} C{Cc}: {us:8.1f} us  {4 * el / us / 1e3:7.0f} GB/s")
            del sets, fa, ba, ba1
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
