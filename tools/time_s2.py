"""Developer timing: stride-2 dgrad (mode 2) at B=512."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from generative_models_b200 import lib  # noqa: E402
from tools.time_conv import ptr, timeit  # noqa: E402

dev = "cuda:0"
lib.load().b200_init(0)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
B = 512
for Ho in (14, 7):
    H = 2 * Ho
    dy = torch.randn(B, Ho, Ho, 128, device=dev)
    w = torch.randn(128, 9, 128, device=dev) * 0.03
    out = torch.empty(B, H, H, 128, device=dev)
    res = torch.randn(B, H, H, 128, device=dev)
    for r in (None, res):
        a = lib.ConvArgs(x=ptr(dy), x2=None, w=ptr(w), out=ptr(out), bias=None, emb=None, residual=ptr(r), N=B, H=Ho, W=Ho,
                         C1=128, C2=0, Ho=H, Wo=H, R=3, S=3, stride=2, pad=1, mode=2, w_cout=128, w_cin=128, n_off=0,
                         out_ld=128, emb_ld=0, res_ld=128, x_ld=0, x2_ld=0, round_out=1)
        us = timeit(lambda: lib.call("b200_conv_igemm", a, st)) * 1e3
        print(f"mode2 dY {Ho}x{Ho} -> dX {H}x{H} res={r is not None}: {us:8.1f} us", flush=True)
