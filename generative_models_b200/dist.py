"""Multi-GPU plumbing: one process per GPU, NCCL over NVLink 5 / NVSwitch.

The reference is single-device (gms/main.py:24,78); data parallelism is an addition with exactly one collective on
the training path — `all_reduce(sum)` of the flat fp32 gradient buffer (6 033 665 live floats, one bucket), followed
by the fused Adam with `grad_scale = 1/world` — and none on the sampling path (independent shards by batch).
Everything here is device-agnostic torch.distributed code, so the world_size-2 `gloo` tests exercise it on CPU.
"""
import os

import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """torchrun contract: RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT.  Returns (rank, world, device)."""
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
    else:
        device = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        backend = backend or ("nccl" if device.type == "cuda" else "gloo")
        kw = {"device_id": device} if backend == "nccl" else {}
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, device


def rank():
    return dist.get_rank() if dist.is_initialized() else 0


def shard_range(n, rank, world):
    """[lo, hi) of `n` items for `rank`: contiguous, disjoint, covering, sizes differ by at most one."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_flat_grad(flat_grad, async_op=False):
    """The training path's only collective: in-place sum of (a bucket of) the flat gradient buffer over all ranks.
    async_op=True returns the Work handle (NCCL: the collective runs on the process group's own stream, ordered
    after everything enqueued so far on the current stream; `.wait()` makes the current stream wait for it), so a
    bucket can be reduced while later kernels of the backward still run.  Returns None at world size 1.
    The optimiser scales the summed gradient by 1/world (folded into the fused Adam)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return None
    work = dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, async_op=async_op)
    return work if async_op else None


def broadcast_params(flat_params, src=0):
    """Make every replica start from rank `src`'s weights (one broadcast of the flat parameter buffer)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(flat_params, src=src)


def sample_sharded(model, n, y=None, gather=False):
    """Large-scale sampling (BASELINE.json configs[3]): rank r draws its shard of the n images with no
    communication; `gather=True` concatenates the shards on every rank (one all_gather at the very end)."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    lo, hi = shard_range(n, rank, world)
    ys = y[lo:hi] if y is not None else None
    out = model.sample(hi - lo, ys)
    if not gather or world == 1:
        return out
    sizes = [shard_range(n, r, world) for r in range(world)]
    bufs = [torch.empty((b - a,) + tuple(out.shape[1:]), dtype=out.dtype, device=out.device) for a, b in sizes]
    dist.all_gather(bufs, out)
    return torch.cat(bufs, 0)
