"""world_size-2 `gloo` test of the data-parallel host logic on CPU: batch sharding, the single flat-gradient
allreduce and the 1/world scaling reproduce the global-batch gradient."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from generative_models_b200 import dist as D


def test_shard_range_partitions():
    for n in (0, 1, 7, 8, 25, 65536):
        for world in (1, 2, 3, 8):
            spans = [D.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from torch import nn
    from generative_models_b200.unet_spec import ParamStore, build_module_tree
    from oracle import ddpm_oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.set_num_threads(2)
    r, w, dev = D.init_from_env(backend="gloo")
    assert (r, w, dev.type) == (rank, world, "cpu")
    # replicas start from rank 0's weights
    st = ParamStore(128, "cpu")
    net = build_module_tree(st, nn.Module())
    net.load_state_dict(O.random_state_dict(1 + rank))  # deliberately different, then broadcast
    D.broadcast_params(st.flat, src=0)
    sd0 = O.random_state_dict(1)
    assert torch.equal(st.view("turn.in_layers.2.weight"), sd0["turn.in_layers.2.weight"])
    # global batch of 4, sharded 2 + 2; local loss = mean over the LOCAL batch (grad_scale 1/B_local)
    g = torch.Generator().manual_seed(5)
    B = 4
    x = torch.rand(B, 1, 28, 28, generator=g) * 2 - 1
    y = torch.tensor([1, -1, 7, 3])
    eps = torch.randn(B, 1, 28, 28, generator=g)
    u = torch.rand(B, generator=g) * 0.8 + 0.1
    lo, hi = D.shard_range(B, rank, world)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    loss, _, _ = O.training_loss(sd, x[lo:hi], y[lo:hi], eps[lo:hi], u[lo:hi], "eps")
    loss.backward()
    st.grad.zero_()
    for k, v in sd.items():
        if v.grad is not None:
            st.grad_view(k).copy_(v.grad)
    # the product's form (model.train_step): two buckets — [bucket_split, numel) = out.* / up.* first, asynchronously,
    # then the rest — and the 1/world factor folded into Adam's grad_scale
    assert 0 < st.bucket_split < st.numel
    assert all((st.offsets[k] >= st.bucket_split) == (k.startswith(("up.", "out.")) and ".emb_layers." not in k)
               for k in st.offsets)
    w1 = D.allreduce_flat_grad(st.grad[st.bucket_split:], async_op=True)
    assert D.allreduce_flat_grad(st.grad[:st.bucket_split]) is None
    w1.wait()
    scale = 1.0 / world
    if rank == 0:
        sdg = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
        gl, _, _ = O.training_loss(sdg, x, y, eps, u, "eps")
        gl.backward()
        worst = 0.0
        for k, v in sdg.items():
            if v.grad is None:
                assert st.grad_view(k).abs().max().item() == 0.0
                continue
            got = st.grad_view(k) * scale
            worst = max(worst, ((got - v.grad).abs().max() / v.grad.abs().max().clamp_min(1e-30)).item())
        q.put(worst)
    dist.barrier()
    dist.destroy_process_group()


def test_flat_grad_allreduce_equals_global_batch_gradient():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    worst = q.get(timeout=600)
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    assert worst < 2e-4, worst
