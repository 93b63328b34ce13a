#!/usr/bin/env python
"""bench.py — DDPM hot-path benchmark (BASELINE.json: train img/s and 1000-step sampled img/s on synthetic
1x28x28 MNIST-shaped data).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path (configs[1], the default line)
    python bench.py --impl reference [...]                          # the UNMODIFIED reference on the host CPU cores
    python bench.py --config 2|3|4 [...]                            # the other BASELINE.json configs
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...   # N>1, one rank per GPU

A "step" is one training step (label drop, eps/u draw, q_sample, UNet fwd, loss, UNet bwd, gradient allreduce,
Adam) over one batch of `--batch` images per GPU (weak scaling).  `value` = images/s with the batch resident in HBM;
`e2e` = the same through the public `train_step` with pinned host batches copied in and the loss read back every
step, plus (`e2e.sampling`) the 1000-step sampler through the public `sample(n, y)` with the images read back.
`sampling` times the unguided fused ancestral loop for `--sample-n` images per GPU, sharded with no communication.
`cpu_baseline` / `gpu_torch_baseline`: the unmodified reference (baseline/_ref) on the host cores and in torch eager on
this GPU, measured in the same run.  One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "ddpm_train_images_per_sec"
UNIT = "img/s"
FWD_FLOP_PER_IMG = 4.3902e9
TRAIN_FLOP_PER_IMG = 13.169e9


def env_int(k, d):
    return int(os.environ.get(k, d))


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        if self.idx < 0:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 8] or [r for _, r in self.rows if len(r) >= 8]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[1]) for r in rows)
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": sorted(reasons),
                "power_w_max": max(float(r[3]) for r in rows), "samples": len(rows)}


def synthetic_batch(B, seed, size=28, channels=1):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, channels, size, size, generator=g) * 2 - 1
    y = torch.randint(0, 10, (B,), generator=g)
    return x, y


# ----------------------------------------------------------------------------------------------- reference arm (CPU)
def workload_string(args, precision=None):
    precision = precision or args.precision
    return (f"DDPM UNet C=128 1x28x28, train step, batch {args.batch}/GPU "
            + ("fp32/TF32 (configs[1])" if precision == "tf32" else "bf16 ResBlock convs (configs[2])")
            + f" + {args.sample_steps}-step ancestral sampling of {args.sample_n} img/GPU")


def run_cpu_reference(steps, warmup, bs, sample_n, sample_steps, timesteps, timeout=1500):
    """The UNMODIFIED reference (baseline/_ref, installed by tools/vendor_reference.sh) through gms.main's own
    construction path, on the host cores, in a child process that cannot see the GPU (so torch.cuda.amp is inert, as on
    a CPU-only machine): oracle/cpu_reference_bench.py.  Returns its JSON, or {"error": ...}."""
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "cpu_reference_bench.py"), "--steps", str(steps), "--warmup",
           str(warmup), "--bs", str(bs), "--sample-n", str(sample_n), "--sample-steps", str(sample_steps),
           "--timesteps", str(timesteps)]
    try:
        r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout)
        lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        if r.returncode != 0 or not lines:
            return {"error": (r.stderr or r.stdout)[-400:]}
        return json.loads(lines[-1])
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)[:400]}


def cpu_sample_string(res, bs):
    t, s = res["train"], res.get("sampling")
    what = ("unmodified reference (gms.main load_model_and_data + the train loop's inner body, main.py:205-215)"
            if res.get("kind") != "port" else
            "oracle port of the reference (oracle/ddpm_oracle.py + autograd + torch.optim.Adam; baseline/_ref absent on this box)")
    out = (f"{what} on CPU, "
           f"{t['steps']} train_step calls at batch {bs} (BASELINE.json configs[0]), {res['threads']} torch threads, "
           f"{res.get('cpu', '')}, torch {res.get('torch', '')}")
    if s:
        out += (f"; sampling: {s['measured_steps']} ancestral steps of {s['n']} images extrapolated to "
                f"{s['extrapolated_to_steps']}")
    return out


def run_reference(args, rank):
    """`--impl reference`: the reference's own CPU implementation of the path on the box's host cores, on configs[0]'s
    batch (64) as a bounded sample of this arm's workload."""
    if rank != 0:
        return 0
    res = run_cpu_reference(max(1, args.steps), max(1, min(args.warmup, 5)), args.cpu_batch, 16, args.cpu_sample_steps,
                            args.sample_steps)
    if "error" in res:
        print(json.dumps({"impl": "reference", "unavailable": "cpu reference failed: " + res["error"][-200:].replace("\n", " ")}),
              file=RESULT_OUT, flush=True)
        return 0
    rate, sps = res["train"]["img_s"], res["train"]["s_per_step"]
    smp = res.get("sampling") or {}
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args), "mean_type": "eps", "sampler": "noisy", "timesteps": args.sample_steps,
                   "bounded_sample": f"batch {args.cpu_batch} per step on the host CPU (configs[0]'s batch) instead of {args.batch}"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": res["threads"], "kind": res.get("kind", "reference"),
                         "sample": cpu_sample_string(res, args.cpu_batch)},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "sampling": {"unguided_img_s": (smp.get("unguided") or {}).get("img_s_1000_steps"),
                             "guided_img_s": (smp.get("guided") or {}).get("img_s_1000_steps"), "n": smp.get("n"),
                             "T": smp.get("extrapolated_to_steps"), "measured_steps": smp.get("measured_steps")}},
        "last_loss": res["train"]["last_loss"],
    }
    print(json.dumps(line), file=RESULT_OUT, flush=True)
    return 0


# ----------------------------------------------------------------------------------------------- this repo's arm (GPU)
def conv_flops(op):
    """Algorithmic FLOPs of one tensor-core launch (2*M*N*K of the implicit GEMM)."""
    a = op.args
    if op.name == "b200_conv_igemm":
        M = a.N * a.Ho * a.Wo
        K = a.R * a.S * (a.C1 + a.C2)
        return 2.0 * M * 128 * K * op.alg_scale  # alg_scale = 1/4 for the zero-stuffed stride-2 dgrads
    if op.name == "b200_conv_wgrad":
        M = a.N * a.Ho * a.Wo
        return 2.0 * M * 128 * a.R * a.S * (a.C1 + a.C2)
    return 0.0


def profile_ops(model, bound, kind):
    """Per-op device time with CUDA events on the launching stream, one eager pass (not a graph replay)."""
    groups = []
    p = bound.plan
    if kind == "train":
        groups = [bound.zero_ops, bound.pre_ops, p.refresh_ops, p.embed_ops[0], p.body_ops, p.head_ops[0],
                  bound.loss_ops, p.refreshT_ops, p.bwd_ops, bound.adam_ops]
    else:
        groups = [p.embed_ops[0], p.body_ops, p.head_ops[0], bound._step_ops(None)[0]]
    ops = [o for g in groups for o in g]
    st = torch.cuda.current_stream()
    sp = model._stream()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(len(ops) + 1)]
    evs[0].record(st)
    for i, o in enumerate(ops):
        rc = o.fn(o.ref, o.ws, o.wsb, sp)
        assert rc == 0, o.name
        evs[i + 1].record(st)
    torch.cuda.synchronize()
    out = {}
    for i, o in enumerate(ops):
        name = o.name
        if name == "b200_conv_igemm":
            from generative_models_b200 import lib as _lib
            strip = _lib.load().b200_conv_kernel_of(o.ref) == 1
            prec = "bf16" if o.args.op_dtype == 1 else "tf32"
            is_dgrad = o.args.mode == 1 or getattr(o, "tag", None) == "dgrad"
            name = (f"conv_strip_kernel<{prec}>" if strip else "conv_igemm_tf32_kernel") + ("[dgrad]" if is_dgrad else "[fprop]")
        elif name == "b200_conv_wgrad":
            name = f"conv_wgrad_kernel<{'bf16' if o.args.op_dtype == 1 else 'tf32'}>"
        d = out.setdefault(name, {"ms": 0.0, "launches": 0, "flops": 0.0})
        d["ms"] += evs[i].elapsed_time(evs[i + 1])
        d["launches"] += 1
        d["flops"] += conv_flops(o)
    return out


def measure_tf32_peak():
    """cuBLAS TF32 8192^3 matmul, best of 5 (the TF32 counterpart of MEASURED_PEAKS.json's bf16 burst figure)."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(8192, 8192, device="cuda")
        b = torch.randn(8192, 8192, device="cuda")
        for _ in range(2):
            a @ b
        best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2 * 8192 ** 3 / best / 1e9
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


class _NullWriter:
    """SummaryWriter stand-in for timing evaluate() (tensorboard's video encoder is not what is measured)."""

    def add_image(self, *a, **k):
        pass

    def add_video(self, *a, **k):
        pass


def run_gpu_torch_baseline(batch, variants="ref_amp_fp16,tf32,bf16", steps=5, sample_steps=5):
    """tools/gpu_torch_baseline.py in a child process: the unmodified reference in torch eager on this GPU."""
    cmd = [sys.executable, os.path.join(ROOT, "tools", "gpu_torch_baseline.py"), "--batch", str(batch), "--steps", str(steps),
           "--warmup", "2", "--sample-n", str(batch), "--sample-steps", str(sample_steps), "--variants", variants]
    env = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    try:
        r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
        lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        if r.returncode != 0 or not lines:
            return {"error": (r.stderr or r.stdout)[-300:]}
        return json.loads(lines[-1])
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)[:300]}


def run_config3(args, rank, world, local_rank):
    """BASELINE.json configs[3]: 65 536 images x 1000 ancestral steps, sharded by batch over the ranks, no communication
    (dist.sample_sharded; every rank walks its shard in chunks of --sample-chunk images so the activations of one chunk
    are what is resident).  value = whole-job sampled images/s, device time, max over ranks."""
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    import __graft_entry__ as ge
    if local_rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    from generative_models_b200 import dist as bdist
    from generative_models_b200 import lib
    from generative_models_b200.model import DiffusionModelB200, default_G
    torch.manual_seed(0)
    T, total = args.sample_steps, args.config3_images
    m = DiffusionModelB200(default_G(device=dev, mean_type="eps", sampler="noisy", timesteps=T, graph=1, precision=args.precision))
    gen = torch.Generator().manual_seed(1)
    for k, p in m.net.named_parameters():
        if ".out_layers.3." in k:
            p.data.copy_((torch.randn(p.shape, generator=gen) * 0.02).to(dev))
    lo, hi = bdist.shard_range(total, rank, world)
    chunk = min(args.sample_chunk, hi - lo)
    clocks = ClockSampler(local_rank if rank == 0 else -1)
    clocks.start()
    b = m._plan(chunk, "sample")  # warm-up: eager, capture, replay
    b.begin_sampling(torch.randn(chunk, 1, 28, 28, device=dev), torch.zeros(chunk, dtype=torch.long, device=dev), None)
    for _ in range(3):
        b.run_graphed(b.enqueue_sampler_step, ("sampler_step", None))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    done, finite = lo, True
    while done < hi:
        n_ = min(chunk, hi - done)
        n_pad = chunk  # a ragged last chunk is padded to the captured plan's size; its first n_ images are kept
        labels = (torch.arange(done, done + n_pad, device=dev) % 10)
        init = torch.randn(n_pad, 1, 28, 28, device=dev)
        z = m._sample_loop(init, labels, None)[0][:n_]
        finite = finite and bool(torch.isfinite(z).all().item())
        done += n_
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t1 = time.time()
    clocks.stop()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
    if rank == 0:
        line = {"metric": "ddpm_sampled_images_per_sec", "value": total / (ms / 1e3), "unit": UNIT, "n_gpus": world,
                "steps": T, "warmup": 3, "ms_per_step": ms / (T * ((hi - lo + chunk - 1) // chunk)), "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
                "config": {"workload": f"configs[3]: {total} images x {T} ancestral steps (unguided), sharded by batch over "
                                       f"{world} GPU(s), chunks of {chunk} images, no communication",
                           "images_per_gpu": hi - lo, "parallelism": f"dp{world} (independent shards)", "cuda_graph": True},
                "seconds": ms / 1e3, "wall_seconds": t1 - t0, "finite": finite,
                "tensor_tflops_per_gpu": (hi - lo) * T * FWD_FLOP_PER_IMG / (ms / 1e3) / 1e12,
                "gpu_launches": None, "clocks": clocks.summary(t0, t1)}
        print(json.dumps(line), file=RESULT_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    import __graft_entry__ as ge
    if local_rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    from generative_models_b200 import dist as bdist
    from generative_models_b200 import lib
    from generative_models_b200.model import DiffusionModelB200, default_G

    B = args.batch
    torch.manual_seed(0)

    def make_model(precision):
        torch.manual_seed(0)
        G = default_G(device=dev, mean_type="eps", sampler="noisy", timesteps=args.sample_steps, bs=B, graph=1,
                      precision=precision, img_channels=args.img_channels, image_size=args.image_size, attention=args.attention)
        m = DiffusionModelB200(G)
        # reference init at seed 0, then a small fill of the zero-initialised convs so every layer does real work
        gen = torch.Generator().manual_seed(1)
        for k, p in m.net.named_parameters():
            if ".out_layers.3." in k:
                p.data.copy_((torch.randn(p.shape, generator=gen) * 0.02).to(dev))
        return m

    model = make_model(args.precision)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def over_ranks(ms, op):
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=op)
            return t.item()
        return ms

    def max_over_ranks(ms):
        return over_ranks(ms, dist.ReduceOp.MAX)

    # ---- device-resident leg: batches already in HBM
    nb = 4
    S_img, C_img = (args.image_size or 28), args.img_channels
    flop_scale = (S_img / 28.0) ** 2  # conv FLOPs scale with the pixel count (the two edge convs also with C_img: < 0.1 %)
    host = [synthetic_batch(B, 100 * rank + i, S_img, C_img) for i in range(nb)]
    host = [(x.pin_memory(), y.pin_memory()) for x, y in host]
    devb = [(x.to(dev), y.to(dev)) for x, y in host]
    W, K = max(args.warmup, 3), args.steps
    # nvidia-smi attaches to every GPU when it starts, which stalls kernel launches on all of them for tens of ms: only
    # rank 0 samples (its own GPU), and it starts before the warm-up so that the start-up cost is outside the timed region
    clocks = ClockSampler(local_rank if rank == 0 else -1)
    clocks.start()
    for i in range(max(W, 8)):  # extra untimed steps beyond W: graph capture + clock ramp
        x, y = devb[i % nb]
        model.train_step(x, y.clone())
    torch.cuda.synchronize()
    l0 = lib.launch_count()
    bound = model._plan(B, "train")
    bound.enqueue_fwd_bwd()
    bound.enqueue_adam()
    launches_per_step = lib.launch_count() - l0

    def settle(mdl, seconds=1.5):
        """Untimed steps (on top of the W warm-up steps) until the GPU has been under this load for `seconds`: the
        first ~second after the load sets in is where the power controller overshoots and re-settles the SM clock
        (seen as a 10-15 % slower 20-step window on some boxes when the timed region started right after W steps)."""
        t0, n = time.time(), 0
        while time.time() - t0 < seconds:
            for i in range(10):
                xs, ys = devb[i % nb]
                mdl.train_step(xs, ys.clone())
            torch.cuda.synchronize()
            n += 10
        return n

    settle_steps = settle(model)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.time()
    e0.record()
    for i in range(K):
        x, y = devb[i % nb]
        model.train_step(x, y.clone())
    e1.record()
    barrier()
    t_wall1 = time.time()
    ms_local = e0.elapsed_time(e1)
    ms = max_over_ranks(ms_local)
    ms_min = over_ranks(ms_local, dist.ReduceOp.MIN)
    value = B * world * K / (ms / 1e3)

    # ---- end-to-end leg: pinned host batches in, loss out, every step (what gms/main.py:205-215 does per batch)
    barrier()
    e0.record()
    last = 0.0
    for i in range(K):
        xh, yh = host[i % nb]
        x = xh.to(dev, non_blocking=True)
        y = yh.to(dev, non_blocking=True)
        last = model.train_step(x, y)["loss"].item()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    t_wall2 = time.time()
    clocks.stop()
    e2e_value = B * world * K / (ms_e2e / 1e3)

    # ---- the collective alone (world > 1): the two gradient buckets, back to back, CUDA events, max over ranks
    allreduce_ms = None
    if world > 1:
        st = model.store
        for _ in range(3):
            bdist.allreduce_flat_grad(st.grad[st.bucket_split:])
            bdist.allreduce_flat_grad(st.grad[:st.bucket_split])
        barrier()
        e0.record()
        for _ in range(10):
            bdist.allreduce_flat_grad(st.grad[st.bucket_split:])
            bdist.allreduce_flat_grad(st.grad[:st.bucket_split])
        e1.record()
        barrier()
        allreduce_ms = max_over_ranks(e0.elapsed_time(e1)) / 10

    # ---- sampling legs: fused ancestral loop, sharded by batch, no communication
    n, T = args.sample_n, args.sample_steps

    def warm_sampler(mdl, n_, guided):
        labels = (torch.arange(n_, device=dev) % 10)
        init = torch.randn(n_, C_img, S_img, S_img, device=dev)
        bs = mdl._plan(n_, "sample_g" if guided else "sample")  # warm-up: 3 steps of the real plan (eager, capture, replay)
        bs.begin_sampling(init, labels, torch.ones(n_, device=dev) if guided else None)
        for _ in range(min(3, T)):
            bs.run_graphed(bs.enqueue_sampler_step, ("sampler_step", None))
        return init, labels

    def time_sampling(mdl):
        """Unguided (the evaluate() path: one net eval per step), z_T already on the device."""
        if T <= 0 or n <= 0:
            return None
        init, labels = warm_sampler(mdl, n, False)
        barrier()
        e0.record()
        z = mdl._sample_loop(init, labels, None)[0]
        e1.record()
        barrier()
        ms_s = max_over_ranks(e0.elapsed_time(e1))
        return {"value": n * world / (ms_s / 1e3), "unit": "img/s", "n_per_gpu": n, "timesteps": T, "guided": False,
                "seconds": ms_s / 1e3, "tensor_tflops": n * T * FWD_FLOP_PER_IMG * flop_scale / (ms_s / 1e3) / 1e12,
                "finite": bool(torch.isfinite(z).all().item())}

    def time_sample_api(mdl, n_):
        """The public call gms/main.py:120 makes — model.sample(n, y): classifier-free guidance, two net evaluations
        per step as ONE 2n-row forward — with the D2H read of the result inside the timed region."""
        if T <= 0 or n_ <= 0:
            return None
        _, labels = warm_sampler(mdl, n_, True)
        barrier()
        t0 = time.perf_counter()
        e0.record()
        out = mdl.sample(n_, labels).cpu()
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        ms_s = max_over_ranks(e0.elapsed_time(e1))
        return {"img_s": n_ * world / (ms_s / 1e3), "seconds": ms_s / 1e3, "wall_seconds": wall, "n": n_ * world, "T": T,
                "guided": True, "net_evals_per_step": 2, "ms_per_step": ms_s / T, "d2h_bytes": out.numel() * 4,
                "tensor_tflops": 2 * n_ * T * FWD_FLOP_PER_IMG * flop_scale / (ms_s / 1e3) / 1e12,
                "finite": bool(torch.isfinite(out).all().item())}

    samp = time_sampling(model)
    samp_api = time_sample_api(model, n) if args.sample_api else None

    # small-n sampling (launch-bound): the calls gms.main makes by default — sample(16, y) of configs[0], evaluate()'s 25
    small = None
    if rank == 0 and args.small_n and T > 0:
        small = {}
        sbound = model._plan(16, "sample_g")
        warm_sampler(model, 16, True)
        l0 = lib.launch_count()
        sbound.enqueue_sampler_step()  # one eager step: every kernel of a guided 2n-row step (replays are not counted)
        small["launches_per_guided_step"] = lib.launch_count() - l0 + 1  # + torch's Philox normal_ for the step noise
        r = time_sample_api(model, 16) if world == 1 else None
        if r:
            small["sample16_guided"] = {"ms_per_step": r["ms_per_step"], "seconds": r["seconds"], "wall_seconds": r["wall_seconds"]}
        if world == 1:
            warm_sampler(model, 25, False)
            model.evaluate(_NullWriter(), None, None, 0)  # capture the u8-history graph
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            model.evaluate(_NullWriter(), None, None, 0)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            small["evaluate25"] = {"ms_per_step": dt / T * 1e3, "wall_seconds": dt, "history": "3 x uint8 [T,25,1,28,28] written by the step kernel"}
        del sbound

    # ---- BASELINE.json configs[2]: the same step with bf16 tensor-core operands in the ResBlock convolutions
    # (reported beside the headline, which stays the fp32/TF32 parity configuration)
    bf16 = None
    if args.precision == "tf32" and args.bf16_leg:
        m2 = make_model("bf16")
        for i in range(max(W, 6)):
            x, y = devb[i % nb]
            m2.train_step(x, y.clone())
        settle(m2)
        barrier()
        e0.record()
        for i in range(K):
            x, y = devb[i % nb]
            m2.train_step(x, y.clone())
        e1.record()
        barrier()
        ms_b = max_over_ranks(e0.elapsed_time(e1))
        bound2 = m2._plan(B, "train")
        prof2 = profile_ops(m2, bound2, "train") if rank == 0 else None
        prof2 = profile_ops(m2, bound2, "train") if rank == 0 else None  # second pass: warm
        sb = time_sampling(m2)
        bf16 = {"config": "configs[2]: ResBlock 3x3 convs with bf16 operands / fp32 accumulate, everything else fp32/TF32; "
                          f"batch {B}/GPU",
                "train_img_s": B * world * K / (ms_b / 1e3), "ms_per_step": ms_b / K,
                "tensor_tflops": B * K / (ms_b / 1e3) * TRAIN_FLOP_PER_IMG * flop_scale / 1e12,
                "sampling_img_s": sb["value"] if sb else None, "sampling_seconds": sb["seconds"] if sb else None,
                "activation_bytes": bound2.plan.activation_bytes()}
        if prof2:
            tot2 = sum(d["ms"] for d in prof2.values())
            bms = sum(d["ms"] for k, d in prof2.items() if k.startswith("conv_strip_kernel<bf16>"))
            bfl = sum(d["flops"] for k, d in prof2.items() if k.startswith("conv_strip_kernel<bf16>"))
            bn = sum(d["launches"] for k, d in prof2.items() if k.startswith("conv_strip_kernel<bf16>"))
            try:
                pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
                peak_bf, src = pk.get("bf16_tflops_sustained") or pk.get("bf16_tflops"), "MEASURED_PEAKS.json bf16 sustained (kernel timed inside a long step)"
            except Exception:
                peak_bf, src = 1391.9, "fallback: the pool's sustained cuBLAS bf16 figure"
            if bms > 0 and peak_bf:
                bf16["roofline"] = {"kernel": "conv_strip_kernel<bf16> (all its fprop + dgrad launches of one train step)",
                                    "bound": "tensor", "achieved": bfl / (bms / 1e3) / 1e12, "peak": peak_bf,
                                    "unit": "TFLOP/s", "frac": bfl / (bms / 1e3) / 1e12 / peak_bf, "traffic": None,
                                    "peak_source": src, "launches": bn, "avg_launch_us": bms / bn * 1e3,
                                    "share_of_step": bms / tot2}
            bf16["breakdown"] = {k: {"ms": round(d["ms"], 4), "launches": d["launches"], "share": round(d["ms"] / tot2, 4),
                                     **({"tflops": round(d["flops"] / (d["ms"] / 1e3) / 1e12, 1)} if d["flops"] else {})}
                                 for k, d in sorted(prof2.items(), key=lambda kv: -kv[1]["ms"])[:10]}
        del m2, bound2
        torch.cuda.empty_cache()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel, measured live with CUDA events (eager pass, one event pair per launch)
    prof = profile_ops(model, bound, "train")
    prof = profile_ops(model, bound, "train")  # second pass: warm
    total_ms = sum(d["ms"] for d in prof.values())
    strip_name = f"conv_strip_kernel<{args.precision}>"
    conv_ms = sum(d["ms"] for k, d in prof.items() if k.startswith(strip_name))
    conv_fl = sum(d["flops"] for k, d in prof.items() if k.startswith(strip_name))
    conv_n = sum(d["launches"] for k, d in prof.items() if k.startswith(strip_name))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    tf32_live = measure_tf32_peak()
    achieved = conv_fl / (conv_ms / 1e3) / 1e12
    roofline = {
        "kernel": strip_name + " (all its fprop + dgrad launches of one train step; the top kernel of the ncu launch "
                  "list of this command under profiles/)", "bound": "tensor",
        "achieved": achieved, "peak": tf32_live, "unit": "TFLOP/s", "frac": achieved / tf32_live,
        # dram__bytes_read.sum + dram__bytes_write.sum of one captured launch (28x28 128->128 3x3 fprop, batch 512):
        # profiles/r01e_ncu_full_summary.md (206.5 MB read + 156.5 MB written; the write-back of the tail sits in L2 when
        # the kernel ends); algorithmic bytes of that launch = 205.5 MB in + 205.5 MB out + 0.6 MB weights
        "traffic": 363.0e6, "traffic_launch": "fprop 3x3 128->128 @28x28, N=512 (algorithmic 411.6e6 B)",
        "peak_source": "cuBLAS TF32 8192^3 measured live in this run (MEASURED_PEAKS.json has no TF32 figure; its bf16 "
                       f"burst is {peaks.get('bf16_tflops')} TFLOP/s, i.e. nominal TF32 = half = "
                       f"{(peaks.get('bf16_tflops') or 0) / 2:.1f})",
        "launches": conv_n, "avg_launch_us": conv_ms / conv_n * 1e3, "share_of_step": conv_ms / total_ms,
    }
    breakdown = {k: {"ms": round(d["ms"], 4), "launches": d["launches"], "share": round(d["ms"] / total_ms, 4),
                     **({"tflops": round(d["flops"] / (d["ms"] / 1e3) / 1e12, 1)} if d["flops"] else {})}
                 for k, d in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}
    act_bytes = bound.plan.activation_bytes()
    del model, bound
    torch.cuda.empty_cache()

    # ---- comparison arms measured beside it: the unmodified reference on the host cores (bounded sample) and, on this
    # same GPU, in torch eager (cuDNN / cuBLAS)
    cpu = run_cpu_reference(args.cpu_steps, 1, args.cpu_batch, 16, min(args.cpu_sample_steps, 6), T) if (world == 1 and args.config == 1) else None
    if cpu and "error" not in cpu:
        cpu_baseline = {"value": cpu["train"]["img_s"], "unit": UNIT, "cores": cpu["threads"], "kind": cpu.get("kind", "reference"),
                        "sample": cpu_sample_string(cpu, args.cpu_batch),
                        "sampling_img_s_unguided": ((cpu.get("sampling") or {}).get("unguided") or {}).get("img_s_1000_steps"),
                        "sampling_img_s_guided": ((cpu.get("sampling") or {}).get("guided") or {}).get("img_s_1000_steps")}
    else:
        cpu_baseline = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference",
                        "sample": "not run (only at N=1 of the default config)" if cpu is None else "failed: " + cpu.get("error", "")[-200:]}
    gpu_base = run_gpu_torch_baseline(B) if (world == 1 and args.gpu_baseline) else None

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "settle_steps": settle_steps,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if args.config == 2 else "weak",
        "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": (f"configs[4] (extension: no reference counterpart): DDPM UNet C=128 with self-attention at 16x16 on 3x64x64, batch {B}/GPU "
                                f"{args.precision}, train step + {T}-step sampling of {n} img/GPU") if args.config == 4 else
                   workload_string(args) if args.config != 2 else
                   f"configs[2]: DDPM data-parallel training, bf16 ResBlock convs, GLOBAL batch {B * world} = {B}/GPU x {world}",
                   "global_batch": B * world, "mean_type": "eps", "sampler": "noisy", "timesteps": T,
                   "parallelism": f"dp{world}", "cuda_graph": True,
                   "l2": "per-step working set (~22 MB/img of activations) >> 126 MB L2; no flush needed",
                   "settle": "1.5 s of untimed steps after the W warm-up steps, before each timed region"},
        "clocks": clocks.summary(t_wall0, t_wall2),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * (C_img * S_img * S_img * 4 + 8), "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / K, "last_loss": last,
                # the other half of BASELINE.json's metric, through the public model.sample(n, y) with the result read back
                "sampling": samp_api,
                "sampling_unguided_img_s": samp["value"] if samp else None},
        "gpu_launches": launches_per_step * K,
        "gpu_launches_per_step": launches_per_step,
        "ms_per_step_min_over_ranks": ms_min / K,
        "allreduce_ms": allreduce_ms,
        "tensor_tflops": value * TRAIN_FLOP_PER_IMG * flop_scale / 1e12 / world,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "gpu_torch_baseline": gpu_base,
        "sampling": samp,
        "sampling_small_n": small,
        "bf16": bf16,
        "breakdown": breakdown,
        "activation_bytes": act_bytes,
    }
    print(json.dumps(line), file=RESULT_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


RESULT_OUT = sys.stdout


def keep_stdout_for_the_result():
    """stdout carries the ONE JSON line and nothing else: fd 1 is pointed at stderr for the rest of the process (NCCL
    prints its version banner to stdout from C), and the result goes to a private duplicate of the original stdout."""
    global RESULT_OUT
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    RESULT_OUT = os.fdopen(saved, "w")


def main():
    keep_stdout_for_the_result()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=512, help="images per GPU per train step")
    ap.add_argument("--sample-n", type=int, default=512, help="images per GPU in the sampling leg")
    ap.add_argument("--sample-steps", type=int, default=1000)
    ap.add_argument("--cpu-batch", type=int, default=64, help="CPU reference arm: images per train step (configs[0]: 64)")
    ap.add_argument("--cpu-steps", type=int, default=6, help="cpu_baseline leg of the GPU arm: timed reference train steps")
    ap.add_argument("--cpu-sample-steps", type=int, default=50, help="CPU reference arm: ancestral steps timed (extrapolated to --sample-steps)")
    ap.add_argument("--sample-api", type=int, default=1, help="time model.sample(n, y) (guided, D2H of the result) for e2e.sampling")
    ap.add_argument("--small-n", type=int, default=1, help="time sample(16, y) and evaluate() (n=25) at full T")
    ap.add_argument("--gpu-baseline", type=int, default=1, help="also time the unmodified reference in torch eager on this GPU")
    ap.add_argument("--precision", default="tf32", choices=["tf32", "bf16"],
                    help="tensor-core operand type of the ResBlock convs (tf32 = configs[1], the headline)")
    ap.add_argument("--bf16-leg", type=int, default=1, help="also time the bf16 configuration (configs[2]) beside the headline")
    ap.add_argument("--attention", type=int, default=0, help="1: the self-attention extension at the lowest-resolution level")
    ap.add_argument("--img-channels", type=int, default=1)
    ap.add_argument("--image-size", type=int, default=0, help="0 = 28 (the reference's MNIST shape)")
    ap.add_argument("--config", type=int, default=1, choices=[1, 2, 3, 4],
                    help="BASELINE.json configs[k]: 1 = batch 512/GPU fp32/TF32 (the default bench line); 2 = bf16 data-parallel "
                         "training at GLOBAL batch 4096 (4096/N per GPU, strong scaling); 3 = 65536 images x 1000 steps sharded")
    ap.add_argument("--config3-images", type=int, default=65536)
    ap.add_argument("--sample-chunk", type=int, default=8192, help="configs[3]: images per resident sampling plan")
    args = ap.parse_args()
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        return run_reference(args, rank)
    nworld = world if world > 1 else max(1, args.gpus)
    if args.config == 4:
        # BASELINE.json configs[4]: 3 x 64 x 64, base 128 channels — an extension (the reference UNet is 1-channel, 28 x 28)
        args.img_channels, args.image_size, args.attention = 3, 64, 1
        args.batch = args.batch if args.batch != 512 else 64
        args.sample_n = min(args.sample_n, args.batch)
        args.small_n = args.gpu_baseline = 0
    if args.config == 2:
        args.precision, args.batch = "bf16", 4096 // nworld
        args.bf16_leg = args.small_n = args.gpu_baseline = args.sample_api = 0
        args.sample_n = 0
    if world == 1 and args.gpus > 1:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd, stdout=RESULT_OUT)
    if args.config == 3:
        return run_config3(args, rank, world, local_rank)
    return run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
