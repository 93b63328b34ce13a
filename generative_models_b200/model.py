"""DiffusionModelB200 — the drop-in for gms/diffusion/diffusion_model.py:14-111 (class DiffusionModel).

Same plugin contract as the reference's `common.GM` subclasses: class attribute `DG`, ctor
`Model(G=G)`, `loss(x, y) -> (loss, metrics)`, `train_step(x, y) -> metrics`, `sample(n, y)`,
`evaluate(writer, x, y, epoch)`, `save(path)`, and a `state_dict()` with the reference's 160 `net.*` keys.
Everything numeric runs in libb200ddpm.so (hand-written sm_100a kernels) through `plan.UNetPlan`; PyTorch supplies
device memory, the RNG (so that eps / u / per-step noise are the reference's own draws, bit for bit), streams, CUDA
graphs and the NCCL process group.  There is no CPU / eager fallback: constructing the model on a non-CUDA device
raises.
"""
import ctypes as C
import random
from pathlib import Path

import torch
from torch import nn

from . import dist, lib
from .plan import Op, UNetPlan, _p, run_ops
from .schedule import logsnr_schedule_cosine, sampler_table
from .unet_spec import ParamStore, build_module_tree

MEAN_TYPES = {"eps": 0, "x": 1, "v": 2}
SAMPLERS = {"ddim": 0, "noisy": 1, "teacher_test": 0}  # value = the step kernel's update rule (0 ddim, 1 ancestral)


class AttrDict(dict):
    """Same behaviour as gms/common.py:24-26."""
    __setattr__ = dict.__setitem__
    __getattr__ = dict.__getitem__


def default_G(**over):
    """The keys gms/main.py:20-40 puts in G that this model reads, with main.py's defaults."""
    G = AttrDict(lr=3e-4, bs=64, device="cuda", pad32=0, weights_from=Path("."), logdir=Path("logs/"))
    G.update(DiffusionModelB200.DG)
    G.update(over)
    return G


class DiffusionModelB200(nn.Module):
    # every key of the reference's DG (diffusion_model.py:15-29), same python types (they become argparse types)
    DG = AttrDict()
    DG.binarize = 0
    DG.timesteps = 250
    DG.hidden_size = 128
    DG.dropout = 0.0
    DG.sampler = "ddim"
    DG.mean_type = "v"
    DG.eval_heavy = 1
    DG.class_cond = 1
    DG.sample_cond_w = -1.0
    DG.cf_drop_prob = 0.1
    DG.teacher_path = Path(".")
    DG.teacher_mode = "step1"
    DG.lr_scheduler = "none"
    # additions (defaults reproduce the reference's behaviour)
    DG.graph = 1  # replay the train step / sampler step as CUDA graphs
    # 'tf32': fp32 storage, TF32 tensor-core operands everywhere (BASELINE.json configs[1], the parity configuration).
    # 'bf16': the ResBlock 3x3 convolutions (87 % of the FLOPs) take bf16 operands with fp32 accumulation — the
    # counterpart of the reference's torch.autocast step (diffusion_model.py:66) — everything else as in 'tf32'.
    DG.precision = "tf32"
    # 1: draw the classifier-free label-drop mask on the device inside the captured step (no CPU RNG, no mask H2D;
    # the caller's y is left untouched).  0 (default) mirrors diffusion_model.py:67 bit for bit: CPU generator draw,
    # in-place edit of the caller's y.
    DG.device_label_drop = 0
    # image channels: 1 = the reference's MNIST UNet (simple_unet.py:92-94, :41); 3 = the 3x64x64 extension of
    # BASELINE.json configs[4] (together with G.image_size), for which the reference has no counterpart
    DG.img_channels = 1
    DG.image_size = 0  # 0: 28 (or 32 with pad32), as the reference; else the image side (a multiple of 4, <= 64)
    # 1: single-head self-attention after the turn block, at the lowest-resolution level (7x7 tokens at 28x28, 16x16 at
    # 64x64) on the tensor cores — an extension named by the north star / configs[4]; adds 10 `net.attn.*` tensors
    DG.attention = 0

    def __init__(self, G):
        super().__init__()
        self.G = G
        dev = torch.device(G.device)
        if dev.type != "cuda":
            raise lib.B200Error("DiffusionModelB200 runs on a CUDA (sm_100a) device only; there is no CPU fallback")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        if G.mean_type not in MEAN_TYPES:
            raise NotImplementedError(f"mean_type {G.mean_type!r} (the reference's 'both' is broken upstream)")
        if G.sampler not in SAMPLERS:
            raise NotImplementedError(f"sampler {G.sampler!r}")
        if G.get("precision", "tf32") not in ("tf32", "bf16"):
            raise NotImplementedError(f"precision {G.precision!r} (tf32 | bf16)")
        self.bf16 = G.get("precision", "tf32") == "bf16"
        if not 0.0 <= float(G.dropout) < 1.0:
            raise ValueError(f"dropout {G.dropout!r} must be in [0, 1)")
        # nn.Dropout(p) of ResBlock.out_layers (simple_unet.py:171), fused into the GroupNorm kernels; the masks are a
        # hash of (this seed, layer, optimiser step, position) — another stream than torch's Philox, same distribution.
        # Ranks of a data-parallel job draw different masks.
        self.drop_seed = (torch.initial_seed() + 0x9E3779B9 * (1 + dist.rank())) & 0x7FFFFFFF
        if G.get("teacher_mode", "step1") not in ("step1", "step2"):
            raise NotImplementedError(f"teacher_mode {G.teacher_mode!r} (step1 | step2)")
        if torch.cuda.current_device() != dev.index:
            raise lib.B200Error(f"one process per GPU: make {dev} the current device (torch.cuda.set_device) before "
                                "constructing DiffusionModelB200")
        rc = lib.load().b200_init(dev.index)
        if rc != 0:
            raise lib.B200Error(f"b200_init failed: {lib.load().b200_last_error().decode()}")
        self.size = int(G.get("image_size", 0)) or (32 if G.pad32 else 28)
        self.img_channels = int(G.get("img_channels", 1))
        self.attention = bool(int(G.get("attention", 0)))
        self.store = ParamStore(G.hidden_size, dev, self.img_channels, self.attention)
        self.store.init_like_reference()
        self.net = build_module_tree(self.store, nn.Module())
        self.optimizer = None  # the fused flat Adam below replaces torch.optim.Adam (diffusion_model.py:56)
        self.store.ensure_adam_state()
        self.adam_step = torch.zeros(1, dtype=torch.int32, device=dev)
        # {lr, grad_scale} live in device memory: the captured step graph reads them, so a learning-rate change (or a
        # scheduler) never forces a re-capture
        self.hparams = torch.tensor([float(G.lr), 1.0], dtype=torch.float32, device=dev)
        self._hparams_host = (float(G.lr), 1.0)
        self._loss_scale = torch.tensor(1.0)
        self._plans = {}
        self._tables = {}
        self._train_graphs = {}
        self.use_graph = bool(G.get("graph", 1))
        self.device_label_drop = bool(int(G.get("device_label_drop", 0)))
        # ring of pinned slots for the CPU-drawn label-drop mask (slot reuse waits on the event of its last H2D copy)
        self._drop_slots, self._drop_i = [], 0
        # data-parallel replicas start from rank 0's weights whatever each rank's RNG state was
        dist.broadcast_params(self.store.flat)
        # progressive distillation (diffusion_model.py:34-43): the student starts from the teacher's weights, the teacher is
        # a frozen copy with its own flat store; its parameters appear under `teacher_net.*` like the reference's
        self.teacher_store = None
        if Path(G.get("teacher_path", ".")) != Path(".") and Path(G.get("weights_from", ".")) == Path("."):
            print("Loading teacher model")
            self.load_state_dict(torch.load(G.teacher_path, map_location=dev), strict=False)
            self.teacher_store = ParamStore(G.hidden_size, dev, self.img_channels, self.attention)
            self.teacher_store.flat.copy_(self.store.flat)
            self.teacher_net = build_module_tree(self.teacher_store, nn.Module())
            for p in self.teacher_net.parameters():
                p.requires_grad = False
                p.grad = None
        elif G.sampler == "teacher_test":
            raise lib.B200Error("sampler='teacher_test' needs a teacher (teacher_path)")

    def load_state_dict(self, state_dict, strict=True, assign=False):
        out = super().load_state_dict(state_dict, strict=strict)
        dist.broadcast_params(self.store.flat)
        return out

    def set_hparams(self, lr, grad_scale):
        """lr / gradient scale of the fused Adam; written to the device only when they change."""
        if (lr, grad_scale) != self._hparams_host:
            self._hparams_host = (lr, grad_scale)
            self.hparams.copy_(torch.tensor([lr, grad_scale], dtype=torch.float32), non_blocking=False)

    def _drop_mask_to_device(self, drop, device):
        """Asynchronous H2D of the CPU-drawn label-drop mask through a small ring of pinned slots."""
        n = drop.shape[0]
        if not self._drop_slots or self._drop_slots[0][0].shape[0] != n:
            self._drop_slots = [(torch.empty(n, dtype=torch.bool).pin_memory(), torch.cuda.Event()) for _ in range(4)]
            self._drop_used = [False] * 4
        i = self._drop_i = (self._drop_i + 1) % 4
        buf, ev = self._drop_slots[i]
        if self._drop_used[i]:
            ev.synchronize()
        buf.copy_(drop)
        d = buf.to(device, non_blocking=True)
        ev.record(torch.cuda.current_stream(self.device))
        self._drop_used[i] = True
        return d

    # ------------------------------------------------------------------ plumbing
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _apply(self, fn, recurse=True):
        """`.to(device)` from gms/main.py:78 must not re-allocate: parameters are views of the flat store."""
        probe = fn(torch.empty(0, device=self.device))
        if probe.device != self.device or probe.dtype != torch.float32:
            raise lib.B200Error("DiffusionModelB200 parameters live in one flat fp32 CUDA buffer; cannot move/cast them")
        return self

    def save(self, path, test_x=None, test_y=None):
        """gms/common.py:154-156."""
        torch.save(self.state_dict(), Path(path) / "model.pt")

    def _table(self, T):
        if T not in self._tables:
            self._tables[T] = sampler_table(T, self.device)
        return self._tables[T]

    def _plan(self, B, kind):
        """kind: 'train' | 'loss' | 'sample' | 'sample_g' (guided: 2B rows).  At most two plans per kind are kept
        (activations are large)."""
        # dropout follows nn.Module's train() / eval() switch, like the reference's nn.Dropout
        drop = float(self.G.dropout) if (self.training and kind in ("train", "loss")) else 0.0
        key = (kind, B, self.size, int(self.G.timesteps), drop)
        p = self._plans.get(key)
        if p is not None:
            return p
        same = [k for k in self._plans if k[0] == kind]
        if len(same) >= 2:
            old = same[0]
            del self._plans[old]
            self._train_graphs.pop(old, None)
            torch.cuda.empty_cache()
        p = _Bound(self, B, kind, dropout=drop)
        self._plans[key] = p
        return p

    # ------------------------------------------------------------------ training
    def loss(self, x, y, eps=None, u=None, cond_w=None, idx=None):
        """diffusion_model.py:76-80 → gaussian_diffusion.py:81-100,156-172.  Forward only (the reference's callers
        use it under no_grad, main.py:163-176); the optimiser path is train_step.  `eps` / `u` (/ `cond_w`, `idx` with a
        teacher) inject the random draws (tests); by default they are drawn like the reference: randn(x.shape), then
        rand(B) (step2 distillation: randint(T, (B,))), then with a teacher 4*rand(B)."""
        b = self._plan(x.shape[0], "loss")
        b.stage_inputs(x, y, eps, u, cond_w, idx)
        st = self._stream()
        run_ops(b.pre_ops, st)
        b.enqueue_teacher_targets()
        b.plan.forward(st)
        run_ops(b.loss_ops, st)
        loss = b.loss_mean[0].clone()
        return loss, {"loss": loss}

    def train_step(self, x, y, eps=None, u=None, cond_w=None, idx=None):
        """diffusion_model.py:63-74: class-label drop, loss, backward, Adam.  fp32 storage with TF32 tensor-core
        convolutions needs no loss scaling, so `loss_scale` is reported as 1.
        Everything after the copy of (x, y) into the plan's static buffers — the eps / u draws (torch's own Philox
        stream, captured), the log-SNR schedule, q_sample, UNet forward, loss, backward, Adam — is ONE CUDA graph at
        world size 1; with a process group the flat-gradient allreduce sits between two graphs, its Up-path half
        overlapped with the Down-path backward (`_Bound.enqueue_fwd_bwd`)."""
        b = self._plan(x.shape[0], "train")
        dev_drop = self.device_label_drop and y is not None
        if y is not None and not dev_drop:
            # same CPU-generator draw and in-place edit of the caller's y as diffusion_model.py:67
            drop = torch.rand(y.shape[0]) < self.G.cf_drop_prob
            if y.is_cuda:
                # y[drop] = -1 on a device tensor goes through nonzero(), a stream synchronisation every step; the
                # masked fill of a (pinned, asynchronously copied) mask writes the same elements without one
                y.masked_fill_(self._drop_mask_to_device(drop, y.device), -1)
            else:
                y[drop] = -1
        b.copy_batch(x, y)
        world = torch.distributed.get_world_size() if torch.distributed.is_initialized() else 1
        self.set_hparams(float(self.G.lr), 1.0 / world)
        key = (eps is None, u is None and idx is None, dev_drop, cond_w is None)
        b.inject(eps, u, cond_w, idx)
        graphed = (lambda fn, k: b.run_graphed(fn, k)) if self.use_graph else (lambda fn, k: fn())
        if world == 1:
            graphed(lambda: (b.enqueue_stage(*key), b.enqueue_fwd_bwd(), b.enqueue_adam()), ("step",) + key)
        else:
            # the one collective of the path: sum of the flat fp32 gradient over NVLink, as two
            # buckets.  The Up-path half (complete once the backward has walked the out head and the Up blocks) is
            # reduced on NCCL's stream while the Down-path backward still runs; the rest follows the backward.
            st = self.store
            graphed(lambda: (b.enqueue_stage(*key), b.enqueue_fwd_bwd(part="up")), ("fwd_bwd_up",) + key)
            w1 = dist.allreduce_flat_grad(st.grad[st.bucket_split:], async_op=True)
            graphed(lambda: b.enqueue_fwd_bwd(part="down"), "bwd_down")
            dist.allreduce_flat_grad(st.grad[:st.bucket_split])
            if w1 is not None:
                w1.wait()
            graphed(b.enqueue_adam, "adam")
        return {"loss": b.loss_mean[0].clone(), "loss_scale": self._loss_scale}

    def loss_and_grads(self, x, y, eps=None, u=None, cond_w=None, idx=None):
        """loss + backward without the optimiser step (parity tests: compare store.grad views with autograd)."""
        b = self._plan(x.shape[0], "train")
        b.stage_inputs(x, y, eps, u, cond_w, idx)
        b.enqueue_fwd_bwd()
        return b.loss_mean[0].clone(), b

    # ------------------------------------------------------------------ sampling
    def _sample_loop(self, init_z, y, cond_w, history=None, noises=None, net_cond_w=None):
        """gaussian_diffusion.py:245-296 as T replays of one fused step:
        UNet(z, logsnr_t | y ‖ uncond) → randn → b200_sampler_step (x/eps reconstruction, guidance, posterior update,
        `where(i == 0, x, z_s)`, optional trajectory rows) → device loop index -= 1.
        With guidance the conditional and the unconditional evaluation (:174-178) are ONE 2n-row forward: rows [0, n)
        carry the labels, rows [n, 2n) the same z with guide = -1.
        history: None | "f32" | "u8" — the three [T, n, 1, S, S] stacks (:293-296), written by the step kernel itself."""
        n = init_z.shape[0]
        T = int(self.G.timesteps)
        kind = "sample_g" if cond_w is not None else "sample"
        if self.G.sampler == "teacher_test":
            kind += "_t"      # the teacher's own weights (gaussian_diffusion.py:271-278)
        elif net_cond_w is not None:
            kind += "_w"      # the student conditioned on w through cond_w_embed (:252-255)
        b = self._plan(n, kind)
        hist = b.begin_sampling(init_z, y, cond_w, history, net_cond_w=net_cond_w)
        noisy = SAMPLERS[self.G.sampler] == 1
        for k in range(T):
            if noises is not None and noisy:
                b.noise.copy_(noises[k].reshape(n, -1))
                b.enqueue_sampler_step(draw_noise=False)
            elif self.use_graph:
                b.run_graphed(b.enqueue_sampler_step, ("sampler_step", history))
            else:
                b.enqueue_sampler_step()
        z = b.plan.z_in[:n].view(n, self.img_channels, self.size, self.size).clone()
        return z, hist

    def _sampling_weights(self, n, want_guidance):
        """gaussian_diffusion.py:247-257: (cond_w for classifier-free guidance, net_cond_w the net is conditioned on).
        `want_guidance` = the caller passed a cond_w (DiffusionModel.sample does, evaluate does not)."""
        net_cond_w = 4.0 * torch.rand(n, device=self.device) if want_guidance else None
        if self.teacher_store is not None:
            if self.G.sampler == "teacher_test":  # guided DDIM steps of the teacher, weights = net_cond_w
                return net_cond_w, None
            return None, net_cond_w               # student conditioned on w, no guidance
        scw = float(self.G.sample_cond_w)
        if scw != -1.0:
            return torch.full((n,), scw, device=self.device), None
        return net_cond_w, None

    def sample(self, n, y=None):
        """diffusion_model.py:82-87.  With labels: classifier-free guidance, per-sample w = 4*rand(n) unless
        G.sample_cond_w is set (gaussian_diffusion.py:247-257); a distilled student is instead conditioned on w and
        sampled without guidance.  y=None (a TypeError in the reference) samples unconditionally without guidance."""
        with torch.no_grad():
            noise = torch.randn((n, self.img_channels, self.size, self.size), device=self.device)
            cond_w, net_cond_w = self._sampling_weights(n, True) if y is not None else (None, None)
            z, _ = self._sample_loop(noise, y, cond_w, net_cond_w=net_cond_w)
            return z

    def sample_trajectory(self, init_z, y=None, cond_w=None, noises=None, history="f32", net_cond_w=None):
        """The three [T, n, 1, S, S] histories (z, x, eps) GaussianDiffusion.sample returns (gaussian_diffusion.py:293-296);
        history="u8": already passed through evaluate()'s proc() (uint8)."""
        with torch.no_grad():
            _, hist = self._sample_loop(init_z, y, cond_w, history=history, noises=noises, net_cond_w=net_cond_w)
            return [h.clone() for h in hist]

    def evaluate(self, writer, x, y, epoch):
        """diffusion_model.py:89-111: 25 class-conditional samples from seed 0 (guided only when G.sample_cond_w is
        set, gaussian_diffusion.py:257); grid + three grid-videos.  proc() — ((t+1)*127.5).clamp(0,255).uint8 — is
        fused into the sampler step: the trajectories leave the device as uint8, 1 B/px."""
        torch.manual_seed(0)
        noise = torch.randn((25, self.img_channels, self.size, self.size), device=self.device)
        labels = torch.arange(25, dtype=torch.long, device=self.device) % 10
        cond_w, net_cond_w = self._sampling_weights(25, False)
        with torch.no_grad():
            _, hist = self._sample_loop(noise, labels, cond_w, history="u8", net_cond_w=net_cond_w)
        zs, xs, eps = (h.cpu() for h in hist)
        if self.G.pad32:
            zs, xs, eps = (t[..., 2:-2, 2:-2] for t in (zs, xs, eps))
        _write_grid(writer, "samples", zs[-1], epoch)
        _write_gridvid(writer, "sampling_process", zs, epoch)
        _write_gridvid(writer, "diffusion_model/eps", eps, epoch)
        _write_gridvid(writer, "diffusion_model/x", xs, epoch)
        torch.manual_seed(random.randint(0, 2**32))


def _grid(x):
    """'(n1 n2) c h w -> c (n1 h) (n2 w)' with n1 = n2 = 5 (gms/common.py:177-180) without einops."""
    n, c, h, w = x.shape[-4:]
    lead = x.shape[:-4]
    x = x.reshape(*lead, 5, 5, c, h, w)
    nd = len(lead)
    x = x.permute(*range(nd), nd + 2, nd + 0, nd + 3, nd + 1, nd + 4)
    return x.reshape(*lead, c, 5 * h, 5 * w)


def _write_grid(writer, tag, x, epoch):
    writer.add_image(tag, _grid(x), epoch)


def _write_gridvid(writer, tag, x, epoch):
    """gms/common.py:183-193."""
    T = x.shape[0]
    vid = _grid(x)[None]
    if vid.shape[2] == 1:  # grey -> RGB, as common.write_gridvid does for the reference's 1-channel images
        vid = vid.repeat(1, 1, 3, 1, 1)
    writer.add_video(tag, vid, epoch, fps=min(T // 3, 60))


class _Bound:
    """A UNetPlan bound to the diffusion-level buffers and ops of one use: 'train', 'loss' or 'sample'."""

    def __init__(self, model, B, kind, dropout=0.0):
        self.m = model
        self.B, self.kind = B, kind
        dev = model.device
        S = model.size
        D = model.img_channels * S * S
        G = model.G
        f32 = dict(dtype=torch.float32, device=dev)
        self.graphs = {}
        mt = MEAN_TYPES[G.mean_type]
        if kind.startswith("sample"):
            # kinds: sample | sample_g (guided: 2B rows) [+ _w: net conditioned on w via cond_w_embed | _t: teacher weights]
            T = int(G.timesteps)
            self.guided = kind.startswith("sample_g")
            self.cond_embed = kind.endswith("_w")
            store = model.teacher_store if kind.endswith("_t") else model.store
            rows = 2 * B if self.guided else B  # guided: conditional rows, then the same z with guide = -1
            self.sched = model._table(T)
            self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
            self.plan = UNetPlan(store, rows, S, train=False, sched=self.sched, step_dev=self.step_dev,
                                 bf16=model.bf16, cond_w=self.cond_embed)
            self.noise = torch.zeros(B, D, **f32)
            self.cond_w = torch.zeros(B, **f32)
            self.T = T
            self.hist = {}
            self.history = None
            self.mt = mt
            return
        # ---- train / loss.  With a teacher (progressive distillation, gaussian_diffusion.py:105-154) the student is
        # conditioned on cond_w and the loss targets come from one (step1) or two (step2) DDIM steps of the teacher
        self.distill = model.teacher_store is not None
        self.mode = G.get("teacher_mode", "step1") if self.distill else None
        self.plan = UNetPlan(model.store, B, S, train=(kind == "train"), bf16=model.bf16, cond_w=self.distill,
                             dropout=dropout, drop_step=model.adam_step, drop_seed=model.drop_seed)
        p = self.plan
        self.x = torch.zeros(B, D, **f32)
        self.eps = torch.zeros(B, D, **f32)
        self.u = torch.zeros(B, **f32)
        self.loss_b = torch.zeros(B, **f32)
        self.mse2 = torch.zeros(B, 2, **f32)
        self.loss_mean = torch.zeros(1, **f32)
        qa = lib.QSampleArgs(x=_p(self.x), eps=_p(self.eps), logsnr=_p(p.t), z=_p(p.z_in), B=B, D=D)
        self.pre_ops = [Op("b200_q_sample", qa)]
        tgt_x, tgt_e = self.x, self.eps
        if self.distill:
            tgt_x, tgt_e = self._build_teacher(model, B, S, D, mt, f32)
        la = lib.LossArgs(out=_p(p.out), z=_p(p.z_in), x=_p(tgt_x), eps=_p(tgt_e), logsnr=_p(p.t),
                          loss_b=_p(self.loss_b), mse2=_p(self.mse2), dout=_p(p.dout) if kind == "train" else None,
                          loss_mean=_p(self.loss_mean), B=B, D=D, mean_type=mt, grad_scale=1.0 / B,
                          loss_type=1 if self.mode == "step1" else 0)  # 'snr' weighting for step1 (:40-41)
        self.loss_ops = [Op("b200_loss", la)]
        if kind == "train":
            st = model.store
            self.zero_ops = [Op("b200_zero", lib.ZeroArgs(ptr=_p(st.grad), bytes=st.numel * 4))]
            self.adam_args = lib.AdamArgs(p=_p(st.flat), g=_p(st.grad), m=_p(st.exp_avg), v=_p(st.exp_avg_sq),
                                          n=st.numel, lr=3e-4, beta1=0.9, beta2=0.999, eps=1e-8, grad_scale=1.0,
                                          step_dev=_p(model.adam_step), step=0, hparams_dev=_p(model.hparams))
            self.adam_ops = [Op("b200_int_increment", lib.StepAdvanceArgs(step_dev=_p(model.adam_step))),
                             Op("b200_adam", self.adam_args)]

    # ---- progressive distillation: the teacher's targets
    def _build_teacher(self, model, B, S, D, mt, f32):
        """Frozen-teacher inference plan + the target ops.  step1 (:113-121): ONE classifier-free-guided DDIM step of the
        teacher (2B rows: labelled, then guide = -1; the teacher is not conditioned on w) — targets are its x / eps
        predictions.  step2 (:122-149): two DDIM half steps of the w-conditioned teacher, then the x implied by where
        they land (b200_distill_target)."""
        p = self.plan
        dev = self.x.device
        step1 = self.mode == "step1"
        self.idx = torch.zeros(B, dtype=torch.int64, device=dev)
        self.logsnr_s = torch.zeros(B, **f32)
        self.logsnr_mid = torch.zeros(B, **f32)
        self.x_target = torch.zeros(B, D, **f32)
        self.eps_target = torch.zeros(B, D, **f32)
        self.z_teacher = torch.zeros(B, D, **f32)
        self.x_pred_teacher = torch.zeros(B, D, **f32)
        t = self.tplan = UNetPlan(model.teacher_store, 2 * B if step1 else B, S, train=False, bf16=model.bf16,
                                  cond_w=not step1)
        run_ops(t.refresh_ops, model._stream())  # the teacher is frozen: its tensor-core weight shadows are built once
        common = dict(sched=None, step_dev=None, noise=None, n=B, D=D, step=0, mean_type=mt, sampler=0)
        if step1:
            a = lib.SamplerStepArgs(z=_p(p.z_in), out_c=_p(t.out), out_u=_p(t.out, B * D), cond_w=_p(p.cond_w),
                                    z_next=_p(self.z_teacher), x_hist=_p(self.x_target), e_hist=_p(self.eps_target),
                                    logsnr_t_b=_p(p.t), logsnr_s_b=_p(self.logsnr_s), **common)
            self.t_ops = [[Op("b200_sampler_step", a)]]
        else:
            a1 = lib.SamplerStepArgs(z=_p(p.z_in), out_c=_p(t.out), out_u=None, cond_w=None, z_next=_p(t.z_in),
                                     logsnr_t_b=_p(p.t), logsnr_s_b=_p(self.logsnr_mid), **common)
            a2 = lib.SamplerStepArgs(z=_p(t.z_in), out_c=_p(t.out), out_u=None, cond_w=None, z_next=_p(self.z_teacher),
                                     x_hist=_p(self.x_pred_teacher), logsnr_t_b=_p(self.logsnr_mid),
                                     logsnr_s_b=_p(self.logsnr_s), **common)
            dt = lib.DistillTargetArgs(z_teacher=_p(self.z_teacher), z_t=_p(p.z_in), x_pred_teacher=_p(self.x_pred_teacher),
                                       logsnr_t=_p(p.t), logsnr_s=_p(self.logsnr_s), index=_p(self.idx),
                                       x_target=_p(self.x_target), eps_target=_p(self.eps_target), n=B, D=D)
            self.t_ops = [[Op("b200_sampler_step", a1)], [Op("b200_sampler_step", a2), Op("b200_distill_target", dt)]]
        return self.x_target, self.eps_target

    def enqueue_teacher_targets(self):
        """Teacher forward(s) + target kernels, between q_sample and the student's forward (no-op without a teacher)."""
        if not self.distill:
            return
        st = self.m._stream()
        t, p, B = self.tplan, self.plan, self.B
        if self.mode == "step1":
            t.z_in[:B].copy_(p.z_in)
            t.z_in[B:].copy_(p.z_in)
            t.t[:B].copy_(p.t)
            t.t[B:].copy_(p.t)
            t.guide[:B].copy_(p.guide)
            t.guide[B:].fill_(-1)
            t.forward(st, refresh=False)
            run_ops(self.t_ops[0], st)
        else:
            t.z_in.copy_(p.z_in)
            t.t.copy_(p.t)
            t.guide.copy_(p.guide)
            t.cond_w.copy_(p.cond_w)
            t.forward(st, refresh=False)
            run_ops(self.t_ops[0], st)   # z_mid -> t.z_in
            t.t.copy_(self.logsnr_mid)
            t.forward(st, refresh=False)
            run_ops(self.t_ops[1], st)

    # ---- inputs
    def copy_batch(self, x, y):
        """(x, y) -> the plan's static buffers (the only per-step work outside the captured step besides the replay)."""
        B = self.B
        assert x.shape[0] == B and x.dtype in (torch.float32, torch.float64)
        self.x.copy_(x.reshape(B, -1), non_blocking=True)
        if y is None:
            self.plan.guide.fill_(-1)
        else:
            self.plan.guide.copy_(y, non_blocking=True)

    def enqueue_stage(self, draw_eps=True, draw_u=True, dev_drop=False, draw_w=True):
        """The random draws in the reference's order (gaussian_diffusion.py:83 randn(x.shape) first, :94 rand(B) — or, for
        step2 distillation, :90 randint(T, (B,)) — second, with a teacher :107 4*rand_like(u) third) from the device's
        default generator, then the log-SNR per sample.  Capturable: torch registers the Philox offset with the graph,
        replays continue the generator's stream exactly like eager draws."""
        if dev_drop:
            # classifier-free label drop (diffusion_model.py:67) with a device draw; ahead of eps/u so that those keep
            # their order relative to each other
            self.plan.guide.masked_fill_(torch.rand(self.B, device=self.x.device) < float(self.m.G.cf_drop_prob), -1)
        if draw_eps:
            self.eps.normal_()
        T = int(self.m.G.timesteps)
        if self.mode == "step2":  # discrete time for distillation (:87-91)
            if draw_u:
                self.idx.copy_(torch.randint(T, (self.B,), device=self.x.device))
            self.u.copy_((self.idx + 1).to(torch.float32) / T)
        elif draw_u:
            self.u.uniform_()
        # logsnr via the reference's torch expression (diffusion_utils.py:198-201): bit-identical per sample
        self.plan.t.copy_(logsnr_schedule_cosine(self.u))
        if self.distill:
            if draw_w:
                self.plan.cond_w.copy_(4.0 * torch.rand_like(self.u))
            self.logsnr_s.copy_(logsnr_schedule_cosine(self.u - 1.0 / T))        # :123-124 (NaN for u < 1/T, unused by step1)
            if self.mode == "step2":
                self.logsnr_mid.copy_(logsnr_schedule_cosine(self.u - 0.5 / T))  # :132-133

    def inject(self, eps=None, u=None, cond_w=None, idx=None):
        """Supplied random draws (parity tests) in place of the device generator's."""
        if eps is not None:
            self.eps.copy_(eps.reshape(self.B, -1))
        if u is not None:
            self.u.copy_(u)
        if idx is not None:
            self.idx.copy_(idx)
        if cond_w is not None:
            self.plan.cond_w.copy_(cond_w)

    def stage_inputs(self, x, y, eps=None, u=None, cond_w=None, idx=None):
        """Eager form: copy the batch, inject or draw eps / u (/ idx, cond_w), derive logsnr."""
        self.copy_batch(x, y)
        self.inject(eps, u, cond_w, idx)
        self.enqueue_stage(eps is None, u is None and idx is None, False, cond_w is None)

    # ---- enqueue
    def enqueue_fwd_bwd(self, part=None):
        """part: None = whole step; "up" = everything up to the point where the out/Up gradients are final;
        "down" = the rest of the backward."""
        st = self.m._stream()
        if part in (None, "up"):
            run_ops(self.zero_ops, st)
            run_ops(self.pre_ops, st)
            self.enqueue_teacher_targets()
            self.plan.forward(st)
            run_ops(self.loss_ops, st)
        self.plan.backward(st, part)

    def enqueue_adam(self):
        run_ops(self.adam_ops, self.m._stream())

    def _step_ops(self, history):
        """b200_sampler_step + the loop-index decrement for one history mode (argument structs are built once)."""
        key = ("ops", history)
        if key not in self.hist:
            p, B, D, T = self.plan, self.B, self.plan.Ci * self.plan.S * self.plan.S, self.T
            dev = self.noise.device
            h = None
            if history == "f32":
                h = [torch.empty(T, B, p.Ci, p.S, p.S, dtype=torch.float32, device=dev) for _ in range(3)]
            elif history == "u8":
                h = [torch.empty(T, B, p.Ci, p.S, p.S, dtype=torch.uint8, device=dev) for _ in range(3)]
            elif history == "last":  # single-row x / eps of the most recent step only (teacher-forced parity tests)
                h = [None] + [torch.empty(B, p.Ci, p.S, p.S, dtype=torch.float32, device=dev) for _ in range(2)]
            elif history is not None:
                raise ValueError(history)
            f, u = (h if history in ("f32", "last") else [None] * 3), (h if history == "u8" else [None] * 3)
            a = lib.SamplerStepArgs(z=_p(p.z_in), out_c=_p(p.out), out_u=_p(p.out, B * D) if self.guided else None,
                                    cond_w=_p(self.cond_w) if self.guided else None, noise=_p(self.noise),
                                    sched=_p(self.sched), step_dev=_p(self.step_dev), z_next=_p(p.z_in),
                                    z_dup=_p(p.z_in, B * D) if self.guided else None,
                                    x_hist=_p(f[1]), e_hist=_p(f[2]), z_hist=_p(f[0]), hist_T=T if history in ("f32", "u8") else 0,
                                    z_u8=_p(u[0]), x_u8=_p(u[1]), e_u8=_p(u[2]), n=B, D=D, step=0,
                                    mean_type=self.mt, sampler=SAMPLERS[self.m.G.sampler])
            self.hist[key] = ([Op("b200_sampler_step", a),
                               Op("b200_step_advance", lib.StepAdvanceArgs(step_dev=_p(self.step_dev)))], h)
        return self.hist[key]

    def begin_sampling(self, init_z, y, cond_w, history=None, net_cond_w=None):
        """Load z_T, the labels and the guidance weights; returns the history tensors of this mode (or None).  They
        are owned by the plan (the captured step writes into them) and are overwritten by the next trajectory."""
        B = self.B
        assert (cond_w is not None) == self.guided
        p = self.plan
        p.z_in[:B].copy_(init_z.reshape(B, -1))
        if y is None:
            p.guide.fill_(-1)
        else:
            p.guide[:B].copy_(y)
        if self.guided:
            p.z_in[B:].copy_(p.z_in[:B])
            p.guide[B:].fill_(-1)  # gaussian_diffusion.py:176: guide = -ones
            self.cond_w.copy_(cond_w)
        assert (net_cond_w is not None) == self.cond_embed
        if self.cond_embed:  # the net itself is conditioned on w (distilled student)
            p.cond_w[:B].copy_(net_cond_w)
            if self.guided:
                p.cond_w[B:].copy_(net_cond_w)
        self.history = history
        self.step_ops, hist = self._step_ops(history)
        self.step_dev.fill_(self.T - 1)
        run_ops(p.refresh_ops, self.m._stream())  # weights are fixed for the whole loop
        return hist

    def teacher_step(self, z_t, y, cond_w, index, noise=None, net_cond_w=None):
        """One sampler step at loop index `index` from a supplied z_t (and noise): returns (x_pred, eps_pred, z_next)
        as [n, 1, S, S] tensors.  Parity tests drive every step of a reference trajectory through this."""
        n, S = self.B, self.plan.S
        hist = self.begin_sampling(z_t, y, cond_w, history="last", net_cond_w=net_cond_w)
        self.step_dev.fill_(index)
        if noise is not None:
            self.noise.copy_(noise.reshape(n, -1))
        else:
            self.noise.zero_()
        self.enqueue_sampler_step(draw_noise=False)
        return hist[1].clone(), hist[2].clone(), self.plan.z_in[:n].view(n, self.plan.Ci, S, S).clone()

    def enqueue_sampler_step(self, draw_noise=True):
        st = self.m._stream()
        self.plan.forward(st, refresh=False)  # rows [0, B): conditional, rows [B, 2B) (guided): unconditional
        if draw_noise and SAMPLERS[self.m.G.sampler] == 1:
            self.noise.normal_()  # drawn after the net evals, like gaussian_diffusion.py:241
        run_ops(self.step_ops, st)

    # ---- CUDA graphs
    def run_graphed(self, fn, key):
        g = self.graphs.get(key)
        if g is None:
            if not self.graphs.get(("warm", key)):
                # first call runs eagerly: it is the warm-up (lazy TMA-descriptor encoding, function attributes)
                self.graphs[("warm", key)] = True
                fn()
                return
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn()
            self.graphs[key] = g
        g.replay()
