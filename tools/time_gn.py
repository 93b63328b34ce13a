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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    a = ap.parse_args()
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
