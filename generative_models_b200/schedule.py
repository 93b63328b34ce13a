"""log-SNR schedule of the reference, evaluated with the reference's own torch expressions so that every scalar the
CUDA kernels consume is bit-identical to what `GaussianDiffusion` would have computed on the same device.

Reference: gms/diffusion/diffusion_utils.py:198-201 (_logsnr_schedule_cosine), gaussian_diffusion.py:33-35
(logsnr_min=-20, logsnr_max=20) and :287-290 (the per-step (logsnr_t, logsnr_s) pair of the sampler loop).
"""
import numpy as np
import torch

LOGSNR_MIN, LOGSNR_MAX = -20.0, 20.0
# numpy float64 scalars, exactly as the reference builds them; multiplied into fp32 tensors they act as python
# scalars, so the arithmetic below runs in fp32.
_B = np.arctan(np.exp(-0.5 * LOGSNR_MAX))
_A = np.arctan(np.exp(-0.5 * LOGSNR_MIN)) - _B


def logsnr_schedule_cosine(t):
    """diffusion_utils.py:198-201: -2 log(tan(a t + b))."""
    return -2.0 * torch.log(torch.tan(_A * t + _B))


def sampler_table(num_steps, device):
    """[T, 2] fp32 table, row i = (logsnr_t, logsnr_s) of loop index i, built like gaussian_diffusion.py:288-290
    (int64 index on the target device, `+ 1.0`, `/ num_steps`, then the schedule).  The reference does this on 0-dim
    tensors one step at a time; the same elementwise kernels applied to an arange give the same bits per element
    (tests/test_cpu_host.py::test_sampler_table_matches_scalar_form_and_fixture checks it against the one-at-a-time form)."""
    i = torch.arange(num_steps, device=device)
    lt = logsnr_schedule_cosine((i + 1.0) / num_steps)
    ls = logsnr_schedule_cosine(i / num_steps)
    return torch.stack([lt, ls], dim=1).to(torch.float32).contiguous()


def sampler_table_scalar(num_steps, device):
    """The literal one-step-at-a-time form of gaussian_diffusion.py:287-290 (used by tests to pin sampler_table)."""
    rows = []
    for k in range(num_steps):
        torch_i = torch.tensor(k, device=device)
        lt = logsnr_schedule_cosine((torch_i + 1.0) / num_steps)
        ls = logsnr_schedule_cosine(torch_i / num_steps)
        rows.append(torch.stack([lt, ls]))
    return torch.stack(rows).to(torch.float32).contiguous()
