"""Developer probe for the tcgen05 conv kernels: prints max-norm relative errors vs torch (fp32, TF32 off).

Run on a GPU box:  python tools/dev_conv_check.py
"""
import ctypes as C
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from generative_models_b200 import lib  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def relerr(a, b):
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-20)).item()


def nhwc(t):  # NCHW logical -> NHWC physical contiguous tensor [N,H,W,C]
    return t.permute(0, 2, 3, 1).contiguous()


def w_khwc(w):  # OIHW -> [O][kh][kw][I]
    return w.permute(0, 2, 3, 1).contiguous()


def run_fprop(x, w, stride, pad, x2=None, bias=None, emb=None, residual=None):
    N, C1, H, W = x.shape
    C2 = 0 if x2 is None else x2.shape[1]
    R = w.shape[2]
    Ho = (H + 2 * pad - R) // stride + 1
    Wo = (W + 2 * pad - R) // stride + 1
    xh = nhwc(x)
    x2h = nhwc(x2) if x2 is not None else None
    wk = w_khwc(w)
    out = torch.full((N, Ho, Wo, 128), float("nan"), device=dev)
    resh = nhwc(residual) if residual is not None else None
    a = lib.ConvArgs(x=ptr(xh), x2=ptr(x2h), w=ptr(wk), out=ptr(out), bias=ptr(bias), emb=ptr(emb),
                     residual=ptr(resh), N=N, H=H, W=W, C1=C1, C2=C2, Ho=Ho, Wo=Wo, R=R, S=R, stride=stride, pad=pad,
                     mode=0, w_cout=128, w_cin=C1 + C2, n_off=0, out_ld=128, emb_ld=128, res_ld=128)
    lib.call("b200_conv_igemm", a, stream())
    torch.cuda.synchronize()
    return out.permute(0, 3, 1, 2)


def run_dgrad(dy, w, pad):
    N, Co, H, W = dy.shape
    Cin = w.shape[1]
    R = w.shape[2]
    dyh = nhwc(dy)
    wk = w_khwc(w)
    out = torch.full((N, H, W, Cin), float("nan"), device=dev)
    for n_off in range(0, Cin, 128):
        a = lib.ConvArgs(x=ptr(dyh), x2=None, w=ptr(wk), out=C.c_void_p(out.data_ptr() + 4 * n_off), bias=None,
                         emb=None, residual=None, N=N, H=H, W=W, C1=Co, C2=0, Ho=H, Wo=W, R=R, S=R, stride=1, pad=pad,
                         mode=1, w_cout=Co, w_cin=Cin, n_off=n_off, out_ld=Cin, emb_ld=0, res_ld=0)
        lib.call("b200_conv_igemm", a, stream())
    torch.cuda.synchronize()
    return out.permute(0, 3, 1, 2)


def run_wgrad(x, dy, R, stride, pad, x2=None, k_slices=0):
    N, C1, H, W = x.shape
    C2 = 0 if x2 is None else x2.shape[1]
    Ho, Wo = dy.shape[2], dy.shape[3]
    xh = nhwc(x)
    x2h = nhwc(x2) if x2 is not None else None
    dyh = nhwc(dy)
    dw = torch.zeros((128, R, R, C1 + C2), device=dev)
    a = lib.WgradArgs(x=ptr(xh), x2=ptr(x2h), dy=ptr(dyh), dw=ptr(dw), N=N, H=H, W=W, C1=C1, C2=C2, Ho=Ho, Wo=Wo, R=R,
                      S=R, stride=stride, pad=pad, k_slices=k_slices)
    lib.call("b200_conv_wgrad", a, stream())
    torch.cuda.synchronize()
    return dw.permute(0, 3, 1, 2)


def main():
    lib.load().b200_init(0)
    g = torch.Generator(device=dev).manual_seed(0)
    rn = lambda *s: torch.randn(*s, device=dev, generator=g)
    ok = True

    def report(name, e, tol=2e-3):
        nonlocal ok
        flag = "ok " if e < tol else "BAD"
        if not e < tol:
            ok = False
        print(f"[{flag}] {name}: relerr={e:.3e}", flush=True)

    # ---- probes: single-tap identity weights (exact in TF32 up to input rounding)
    x = rn(8, 128, 28, 28)
    for kh, kw in [(1, 1), (0, 0), (2, 2), (0, 2), (1, 0)]:
        w = torch.zeros(128, 128, 3, 3, device=dev)
        w[:, :, kh, kw] = torch.eye(128, device=dev)
        y = run_fprop(x, w, 1, 1)
        ref = F.conv2d(x, w, padding=1)
        report(f"fprop probe tap({kh},{kw})", relerr(y, ref))
    # permutation weight to check channel order inside swizzled tiles
    perm = torch.randperm(128, device=dev, generator=g)
    w = torch.zeros(128, 128, 3, 3, device=dev)
    w[torch.arange(128, device=dev), perm, 1, 1] = 1.0
    report("fprop probe channel-perm", relerr(run_fprop(x, w, 1, 1), F.conv2d(x, w, padding=1)))

    # ---- random fprop cases
    w = rn(128, 128, 3, 3) * 0.05
    report("fprop 3x3 s1 N8 28x28", relerr(run_fprop(x, w, 1, 1), F.conv2d(x, w, padding=1)))
    x7 = rn(3, 128, 7, 7)
    report("fprop 3x3 s1 N3 7x7 (tail)", relerr(run_fprop(x7, w, 1, 1), F.conv2d(x7, w, padding=1)))
    report("fprop 3x3 s2 28->14", relerr(run_fprop(x, w, 2, 1), F.conv2d(x, w, stride=2, padding=1)))
    x14 = rn(5, 128, 14, 14)
    report("fprop 3x3 s2 14->7", relerr(run_fprop(x14, w, 2, 1), F.conv2d(x14, w, stride=2, padding=1)))
    xb = rn(8, 128, 28, 28)
    w1 = rn(128, 256, 1, 1) * 0.05
    report("fprop 1x1 concat256", relerr(run_fprop(x, w1, 1, 0, x2=xb), F.conv2d(torch.cat([x, xb], 1), w1)))
    w2 = rn(128, 256, 3, 3) * 0.05
    report("fprop 3x3 concat256", relerr(run_fprop(x, w2, 1, 1, x2=xb), F.conv2d(torch.cat([x, xb], 1), w2, padding=1)))
    bias = rn(128)
    emb = rn(8, 128)
    res = rn(8, 128, 28, 28)
    ref = F.conv2d(x, w, bias=bias, padding=1) + emb[:, :, None, None] + res
    report("fprop 3x3 +bias+emb+res", relerr(run_fprop(x, w, 1, 1, bias=bias, emb=emb, residual=res), ref))

    # ---- dgrad
    dy = rn(8, 128, 28, 28)
    ref = torch.nn.grad.conv2d_input(x.shape, w, dy, padding=1)
    report("dgrad 3x3 Cin128", relerr(run_dgrad(dy, w, 1), ref))
    ref = torch.nn.grad.conv2d_input((8, 256, 28, 28), w2, dy, padding=1)
    report("dgrad 3x3 Cin256", relerr(run_dgrad(dy, w2, 1), ref))
    ref = torch.nn.grad.conv2d_input((8, 256, 28, 28), w1, dy, padding=0)
    report("dgrad 1x1 Cin256", relerr(run_dgrad(dy, w1, 0), ref))
    dy7 = rn(3, 128, 7, 7)
    ref = torch.nn.grad.conv2d_input((3, 128, 7, 7), w, dy7, padding=1)
    report("dgrad 3x3 7x7 tail", relerr(run_dgrad(dy7, w, 1), ref))

    # ---- wgrad
    ref = torch.nn.grad.conv2d_weight(x, w.shape, dy, padding=1)
    report("wgrad 3x3 s1", relerr(run_wgrad(x, dy, 3, 1, 1), ref))
    report("wgrad 3x3 s1 1 slice", relerr(run_wgrad(x, dy, 3, 1, 1, k_slices=1), ref))
    dy14 = rn(8, 128, 14, 14)
    ref = torch.nn.grad.conv2d_weight(x, w.shape, dy14, stride=2, padding=1)
    report("wgrad 3x3 s2", relerr(run_wgrad(x, dy14, 3, 2, 1), ref))
    ref = torch.nn.grad.conv2d_weight(torch.cat([x, xb], 1), w1.shape, dy, padding=0)
    report("wgrad 1x1 concat256", relerr(run_wgrad(x, dy, 1, 1, 0, x2=xb), ref))
    ref = torch.nn.grad.conv2d_weight(torch.cat([x, xb], 1), w2.shape, dy, padding=1)
    report("wgrad 3x3 concat256", relerr(run_wgrad(x, dy, 3, 1, 1, x2=xb), ref))
    ref = torch.nn.grad.conv2d_weight(x7, w.shape, dy7, padding=1)
    report("wgrad 3x3 7x7 tail", relerr(run_wgrad(x7, dy7, 3, 1, 1), ref))

    # ---- timing (B=512, 28x28, 128->128 3x3)
    xb512 = nhwc(rn(512, 128, 28, 28))
    wk = w_khwc(w)
    out = torch.empty((512, 28, 28, 128), device=dev)
    a = lib.ConvArgs(x=ptr(xb512), x2=None, w=ptr(wk), out=ptr(out), bias=None, emb=None, residual=None, N=512, H=28,
                     W=28, C1=128, C2=0, Ho=28, Wo=28, R=3, S=3, stride=1, pad=1, mode=0, w_cout=128, w_cin=128, n_off=0,
                     out_ld=128, emb_ld=0, res_ld=0)
    for _ in range(3):
        lib.call("b200_conv_igemm", a, stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        lib.call("b200_conv_igemm", a, stream())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    fl = 2 * 512 * 784 * 128 * 1152
    print(f"fprop B512 3x3 128->128: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
    xt = xb512.permute(0, 3, 1, 2)  # channels_last view
    torch.backends.cudnn.allow_tf32 = True
    for _ in range(3):
        F.conv2d(xt, w, padding=1)
    e0.record()
    for _ in range(10):
        F.conv2d(xt, w, padding=1)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"cudnn TF32 channels_last same conv: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
    print("ALL OK" if ok else "SOME BAD")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
