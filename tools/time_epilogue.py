"""Developer timing: the strip kernel's epilogue variants with bf16 operands (3x3 128->128, B=512)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from generative_models_b200 import lib  # noqa: E402
from tools.time_conv import ptr, timeit  # noqa: E402

dev = "cuda:0"


def main():
    lib.load().b200_init(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    B = 512
    for H in (28, 14):
        xb = torch.randn(B, H, H, 128, device=dev).bfloat16()
        wb = (torch.randn(128, 3, 3, 128, device=dev) * 0.05).bfloat16()
        out = torch.empty(B, H, H, 128, device=dev)
        outb = torch.empty(B, H, H, 128, device=dev, dtype=torch.bfloat16)
        res = torch.randn(B, H, H, 128, device=dev)
        emb = torch.randn(B, 128, device=dev)
        for name, kw in [("fp32 out", {}), ("fp32 out + residual", dict(residual=ptr(res))),
                         ("fp32 out + bf16 copy", dict(out_bf16=ptr(outb))),
                         ("fp32 out + residual + bf16 copy", dict(residual=ptr(res), out_bf16=ptr(outb))),
                         ("bf16 out", dict(out=ptr(outb), out_dtype=1)),
                         ("bf16 out + emb", dict(out=ptr(outb), out_dtype=1, emb=ptr(emb), emb_ld=128))]:
            args = dict(x=ptr(xb), x2=None, w=ptr(wb), out=ptr(out), bias=None, emb=None, residual=None, N=B, H=H, W=H,
                        C1=128, C2=0, Ho=H, Wo=H, R=3, S=3, stride=1, pad=1, mode=0, w_cout=128, w_cin=128, n_off=0,
                        out_ld=128, emb_ld=0, res_ld=128, x_ld=0, x2_ld=0, round_out=0, op_dtype=1)
            args.update(kw)
            a = lib.ConvArgs(**args)
            ms = timeit(lambda: lib.call("b200_conv_igemm", a, st))
            fl = 2.0 * B * H * H * 128 * 9 * 128
            print(f"H={H} {name:34s}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
