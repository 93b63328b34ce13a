"""Helpers for the -m gpu parity tests: raw-pointer plumbing for the C ABI and layout conversions."""
import ctypes as C

import torch


def ptr(t, off=0):
    return C.c_void_p(t.data_ptr() + 4 * off) if t is not None else None


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def rel(a, b):
    """max|a-b| / max|b| — the norm SURVEY.md §8c states for the 1e-3 parity bar."""
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30)).item()


def nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous()


def w_khwc(w):
    return w.permute(0, 2, 3, 1).contiguous()
