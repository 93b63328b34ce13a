"""cuDNN (torch eager, channels_last, cudnn.benchmark) per-layer timings of the Appendix-A conv shapes at batch 512:
fprop, backward-data and backward-weight separately, TF32 and bf16 — the library kernels beside tools/time_conv.py's."""
import json
import os
import sys

import torch

dev = "cuda:0"


def timeit(fn, n=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = True
    B = int(os.environ.get("B", 512))
    res = []
    for (H, Cin, R) in [(28, 128, 3), (28, 256, 3), (28, 256, 1), (14, 128, 3), (14, 256, 3), (7, 128, 3)]:
        for dt in (torch.float32, torch.bfloat16):
            x = torch.randn(B, Cin, H, H, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
            w = (torch.randn(128, Cin, R, R, device=dev, dtype=dt) * 0.05).contiguous(memory_format=torch.channels_last)
            dy = torch.randn(B, 128, H, H, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
            pad = R // 2
            fl = 2.0 * B * H * H * 128 * R * R * Cin
            f = lambda: torch.nn.functional.conv2d(x, w, None, 1, pad)
            bd = lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [1, 1], [pad, pad], [1, 1], False, [0, 0], 1, [True, False, False])
            bw = lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [1, 1], [pad, pad], [1, 1], False, [0, 0], 1, [False, True, False])
            row = {"H": H, "Cin": Cin, "R": R, "dtype": "tf32" if dt == torch.float32 else "bf16"}
            for nm, fn in (("fprop", f), ("dgrad", bd), ("wgrad", bw)):
                ms = timeit(fn)
                row[nm + "_us"] = round(ms * 1e3, 1)
                row[nm + "_tflops"] = round(fl / ms / 1e9, 1)
            res.append(row)
            print(row, flush=True)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
