"""Developer timing of single conv launches through the C ABI (B=512, the Appendix-A layer shapes)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from generative_models_b200 import lib  # noqa: E402

dev = "cuda:0"


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    lib.load().b200_init(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    B = int(os.environ.get("B", 512))
    for (H, Cin, R, mode) in [(28, 128, 3, 0), (28, 256, 3, 0), (28, 256, 1, 0), (14, 128, 3, 0), (7, 128, 3, 0),
                              (28, 128, 3, 1)]:
        C1 = 128
        C2 = Cin - 128 if mode == 0 else 0
        x = torch.randn(B, H, H, 128, device=dev)
        x2 = torch.randn(B, H, H, 128, device=dev) if C2 else None
        w = torch.randn(128, R, R, Cin, device=dev) * 0.05
        out = torch.empty(B, H, H, 128, device=dev)
        res = torch.randn(B, H, H, 128, device=dev)
        eye = torch.eye(128, device=dev)
        xb, x2b, wb = x.bfloat16(), (x2.bfloat16() if x2 is not None else None), w.bfloat16()
        for with_res in (0, 1, 2, 3):
            if with_res == 3:  # bf16 operands, no residual
                a = lib.ConvArgs(x=ptr(xb), x2=ptr(x2b), w=ptr(wb), out=ptr(out), bias=None, emb=None, residual=None, N=B,
                                 H=H, W=H, C1=C1, C2=C2, Ho=H, Wo=H, R=R, S=R, stride=1, pad=R // 2, mode=mode, w_cout=128,
                                 w_cin=Cin, n_off=0, out_ld=128, emb_ld=0, res_ld=128, x_ld=0, x2_ld=0, round_out=1,
                                 op_dtype=1)
                ms = timeit(lambda: lib.call("b200_conv_igemm", a, st))
                fl = 2.0 * B * H * H * 128 * R * R * (Cin if mode == 0 else 128)
                print(f"H={H} Cin={Cin} R={R} mode={mode} bf16: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TFLOP/s", flush=True)
                continue
            a = lib.ConvArgs(x=ptr(x), x2=ptr(x2), w=ptr(w), out=ptr(out), bias=None, emb=None,
                             residual=ptr(res) if with_res == 1 else None, N=B, H=H, W=H, C1=C1, C2=C2, Ho=H, Wo=H, R=R, S=R,
                             stride=1, pad=R // 2, mode=mode, w_cout=128, w_cin=Cin, n_off=0, out_ld=128, emb_ld=0,
                             res_ld=128, x_ld=0, x2_ld=0, round_out=1)
            if with_res == 2:
                a.skip_x, a.skip_w, a.skip_c1 = ptr(res), ptr(eye), 128
            ms = timeit(lambda: lib.call("b200_conv_igemm", a, st))
            fl = 2.0 * B * H * H * 128 * R * R * (Cin if mode == 0 else 128)
            print(f"H={H} Cin={Cin} R={R} mode={mode} res={with_res}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TFLOP/s", flush=True)
    for (H, Cin) in [(28, 128), (28, 256), (14, 128), (7, 128)]:
        x = torch.randn(B, H, H, Cin, device=dev)
        dy = torch.randn(B, H, H, 128, device=dev)
        dw = torch.zeros(128, 3, 3, Cin, device=dev)
        for bfm in (0, 1):
            xx, dd = (x.bfloat16(), dy.bfloat16()) if bfm else (x, dy)
            a = lib.WgradArgs(x=ptr(xx), x2=None, dy=ptr(dd), dw=ptr(dw), N=B, H=H, W=H, C1=Cin, C2=0, Ho=H, Wo=H, R=3, S=3,
                              stride=1, pad=1, k_slices=0, op_dtype=bfm)
            ms = timeit(lambda: lib.call("b200_conv_wgrad", a, st))
            fl = 2.0 * B * H * H * 128 * 9 * Cin
            print(f"wgrad H={H} Cin={Cin} {'bf16' if bfm else 'tf32'}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
