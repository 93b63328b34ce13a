"""Static execution plan of the UNet (gms/diffusion/simple_unet.py:16-186) over libb200ddpm.so.

A plan is built once per (batch, image size, train/infer): it owns every activation buffer (allocated by torch,
handed to the library as raw pointers), every pre-filled C argument struct, and two op lists — forward and, for
training plans, backward.  Running a plan is a flat loop of `fn(&args, ws, ws_bytes, stream)` C-ABI calls: no
autograd graph, no per-step allocation, no host<->device sync, so a whole train step or sampler step can be
captured in a CUDA graph and replayed.

Layout: activations are NHWC fp32 `[B, H*W, C]` pixel rows; the model's 1-channel input/output is `[B, H*W]`
(NCHW == NHWC when C == 1).  `torch.cat([x, skip], 1)` (simple_unet.py:150) is never materialised: every consumer
takes two base pointers.  The reverse pass accumulates parameter gradients straight into the flat gradient buffer of
`unet_spec.ParamStore` (the NCCL allreduce bucket).
"""
import ctypes as C
import os

import zlib

import torch

from . import lib
from .unet_spec import RESBLOCKS

GN_EPS = 1e-5  # nn.GroupNorm default (simple_unet.py:39,162,169)
MAX_TIMESTEPS = 256  # simple_unet.py:13


def _p(t, off_floats=0):
    """raw device pointer of a tensor (+ offset in fp32 elements)"""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr() + 4 * off_floats)


class Op:
    __slots__ = ("fn", "ref", "args", "ws", "wsb", "name", "alg_scale", "tag")

    def __init__(self, name, args, ws=None):
        self.name = name
        self.tag = None  # "dgrad" on backward-data launches (they run the forward kernel on transposed weights)
        self.alg_scale = 1.0  # algorithmic / executed work of this launch (bench.py's roofline accounting)
        self.fn = getattr(lib.load(), name)
        self.args = args  # keep the struct alive
        self.ref = C.byref(args)
        self.ws = C.c_void_p(ws.data_ptr()) if ws is not None else None
        self.wsb = ws.numel() * ws.element_size() if ws is not None else 0


def run_ops(ops, stream_ptr):
    """Enqueue every op on the given raw cudaStream_t; raise on the first rejected call (no fallback)."""
    for op in ops:
        rc = op.fn(op.ref, op.ws, op.wsb, stream_ptr)
        if rc != 0:
            raise lib.B200Error(f"{op.name} failed (rc={rc}): {lib.load().b200_last_error().decode()}")


class _Pool:
    """Buffer pool: training plans keep every tensor (the backward reads them); inference plans recycle."""

    def __init__(self, device, recycle):
        self.device = device
        self.recycle = recycle
        self.free_list = {}
        self.all = []

    def get(self, *shape):
        n = 1
        for s in shape:
            n *= s
        if self.recycle and self.free_list.get(n):
            return self.free_list[n].pop().view(*shape)
        t = torch.empty(n, dtype=torch.float32, device=self.device)
        if os.environ.get("B200_POISON"):  # developer check: a read-before-write shows up as NaN
            t.fill_(float("nan"))
        self.all.append(t)
        return t.view(*shape)

    def release(self, t):
        if self.recycle and t is not None:
            self.free_list.setdefault(t.numel(), []).append(t.view(-1))

    def bytes(self):
        return sum(t.numel() * 4 for t in self.all)


class _Grad:
    """A gradient tensor seen as [B*HW] pixel rows of 128 channels with a pixel stride (a slice of a wider buffer).
    bf: optional dense bf16 copy [B*HW, 128] of the same slice (the dY operand of the bf16 dgrad / wgrad)."""
    __slots__ = ("t", "off", "ld", "bf", "prod")

    def __init__(self, t, off=0, ld=None, bf=None):
        self.t = t
        self.off = off
        self.ld = ld if ld is not None else t.shape[-1]
        self.bf = bf if bf is not None else (getattr(t, "bf16_copy", None) if off == 0 else None)
        # the b200_gn_bwd argument struct that writes t (its first 128 channels = this slice): column sums of the
        # slice are fused into that launch instead of a separate b200_colsum pass
        self.prod = getattr(t, "producer_gn", None) if off == 0 else None

    def ptr(self):
        return _p(self.t, self.off)


class UNetPlan:
    """Forward (and backward) op lists of SimpleUnet for a fixed batch `B` and image side `S`."""

    def __init__(self, store, B, S, train, sched=None, step_dev=None, bf16=False, cond_w=False, dropout=0.0,
                 drop_step=None, drop_seed=0):
        """bf16: the 24 ResBlock 3x3 convolutions (87 % of the FLOPs) take bf16 operands — GroupNorm writes its output
        as bf16, the weights come from a bf16 shadow, the backward's dY operands are bf16 copies written by the
        producing GroupNorm-backward — with fp32 accumulation; the residual stream, GroupNorm statistics, every
        gradient that is summed, and the plain / skip / edge convolutions stay fp32 (TF32 tensor cores)."""
        assert S % 4 == 0, "image side must be divisible by 4 (two stride-2 levels)"
        self.bf16 = bool(bf16)
        # nn.Dropout(p) of ResBlock.out_layers (simple_unet.py:171), fused into that block's second GroupNorm (b200ddpm.h:
        # dropout_p / dropout_seed / dropout_step); drop_step is a device counter that advances once per optimiser step
        self.dropout, self.drop_step, self.drop_seed = float(dropout), drop_step, int(drop_seed)
        # cond_w: the net is conditioned on a per-sample guidance weight through cond_w_embed (simple_unet.py:60-64;
        # progressive distillation: students, and step2 teachers); self.cond_w [B] is the input buffer
        self.use_cond_w = bool(cond_w)
        self.fuse_colsum = True
        # bf16 STORAGE of the tensors only GroupNorm reads (h1, the dgrad outputs), like the reference's autocast conv
        # outputs: halves their HBM traffic (B200_BF16_STORE=0 keeps them fp32, for A/B measurements)
        import os
        self.bfs = self.bf16 and os.environ.get("B200_BF16_STORE", "1") == "1"
        if S * S > 2304:  # the raw-staged GroupNorm backward holds a 16-channel slice (6 B/element) in shared memory
            self.bfs = False
        # dh1 as bf16 only (no fp32 dx) needs the shared-memory-resident GroupNorm backward (16 channels x 8 B/pixel)
        self.dh1_bf_only = self.bf16 and S * S * 128 <= 220 * 1024
        if self.bf16:
            store.ensure_bf16()
        self.store = store
        self.Ci = Ci = getattr(store, "img_channels", 1)  # image channels (1 = the reference; 3 = configs[4] extension)
        self.B, self.S, self.train = B, S, train
        self.Cc = Cc = store.C
        assert Cc == 128, "the tcgen05 conv kernels are specialised for hidden_size=128 (N tile = 128 channels)"
        self.E = E = 2 * Cc
        dev = store.device
        self.pool = _Pool(dev, recycle=not train)
        self.gpool = _Pool(dev, recycle=True)
        f32 = dict(dtype=torch.float32, device=dev)
        # ---- static inputs / outputs
        self.z_in = torch.zeros(B, Ci * S * S, **f32)  # planar [B][Ci][S*S] (NCHW; == NHWC when Ci == 1)
        self.t = torch.zeros(B, **f32)
        self.guide = torch.full((B,), -1, dtype=torch.int64, device=dev)
        self.out = torch.zeros(B, Ci * S * S, **f32)
        self.out2 = torch.zeros(B, Ci * S * S, **f32)
        self.sched, self.step_dev = sched, step_dev
        import math
        # the reference evaluates exp(-math.log(max_period) * arange / half) in fp32 on the CPU (simple_unet.py:216-220)
        self.freqs = torch.exp(-math.log(MAX_TIMESTEPS) * torch.arange(0, 32, dtype=torch.float32) / 32).to(dev)
        # cond_w_embed's timestep_embedding uses max_period = 4 (simple_unet.py:62)
        self.freqs_w = torch.exp(-math.log(4) * torch.arange(0, 32, dtype=torch.float32) / 32).to(dev)
        self.cond_w = torch.zeros(B, **f32)
        self.attention = bool(getattr(store, "attention", False))
        self.n_stats = 25 + (1 if self.attention else 0)
        self.stats = torch.zeros(self.n_stats, B, 64, **f32)
        self._stat_i = 0
        self.records = {}
        self.skips = {}
        self._build_forward()
        if train:
            self.dout = torch.zeros(B, self.Ci * S * S, **f32)
            self._build_backward()

    # ------------------------------------------------------------------ small constructors
    def _w(self, name):
        return C.c_void_p(self.store.ptr(name))

    def _wtc(self, name):
        """tensor-core operand: the TF32-rounded shadow of the parameter (refreshed by self.refresh_ops)"""
        return C.c_void_p(self.store.tc_ptr(name))

    def _wbf(self, name):
        return C.c_void_p(self.store.bf_ptr(name))

    def _g(self, name):
        return C.c_void_p(self.store.gptr(name))

    @staticmethod
    def _get_bf(pool, *shape):
        """storage for a bf16 tensor of `shape` from an fp32 pool (half the floats); only its pointer is used"""
        assert shape[-1] % 2 == 0
        return pool.get(*shape[:-1], shape[-1] // 2)

    def _next_stats(self):
        s = self.stats[self._stat_i]
        self._stat_i += 1
        return s

    def _conv(self, x, x2, wname, out, H, Ho, R, stride, pad, C1, C2, bias=True, emb=None, residual=None,
              res_ld=128, x_ld=0, x2_ld=0, round_out=1, bf=False, out_bf=None, out_dtype=0):
        if out_bf is not None:
            out.bf16_copy = out_bf
        a = lib.ConvArgs(x=_p(x), x2=_p(x2), w=(self._wbf if bf else self._wtc)(wname + ".weight"), out=_p(out),
                         bias=self._w(wname + ".bias") if bias else None, emb=emb,
                         residual=residual, N=self.B, H=H, W=H, C1=C1, C2=C2, Ho=Ho, Wo=Ho, R=R, S=R, stride=stride,
                         pad=pad, mode=0, w_cout=128, w_cin=C1 + C2, n_off=0, out_ld=128, emb_ld=12 * self.Cc,
                         res_ld=res_ld, x_ld=x_ld, x2_ld=x2_ld, round_out=round_out, op_dtype=1 if bf else 0,
                         out_bf16=_p(out_bf), out_dtype=out_dtype)
        return Op("b200_conv_igemm", a)

    def _rel(self, t):
        """release an activation and its bf16 copy (inference plans recycle; training plans keep everything)"""
        if t is not None:
            self.pool.release(t)
            self.pool.release(getattr(t, "bf16_copy", None))

    def _cast_copy(self, ops, t, rows):
        """bf16 mode: dense bf16 copy of a [rows, 128] fp32 activation whose producer cannot dual-write"""
        t.bf16_copy = self._get_bf(self.pool, rows, 128)
        ops.append(Op("b200_cast_bf16", lib.CastArgs(src=_p(t), dst=_p(t.bf16_copy), rows=rows, cols=128, src_ld=0)))

    def _dgrad(self, dy, wname, out, H, R, pad, w_cin, n_off, out_ld, residual=None, res_ld=0, bf=False, out_dtype=0):
        """dy: _Grad with 128 channels; writes 128 input channels [n_off, n_off+128) of `out`.
        bf: read the dense bf16 copy dy.bf and the bf16 weight shadow.
        Backward-data of a stride-1 'same' convolution is the forward convolution of dY with the transposed, tap-flipped
        weights: with the [Cin][T][128] shadow (ParamStore.ensure_transposed) it runs the K-major forward path, which is
        ~12 % faster than the MN-major weight operand of mode 1 (239 -> 211 us at 28x28, 128 channels)."""
        if self.dgrad_T:
            esz = 2 if bf else 4
            wT = (self.store.bfT_ptr if bf else self.store.tcT_ptr)(wname + ".weight") + esz * n_off * R * R * 128
            a = lib.ConvArgs(x=_p(dy.bf) if bf else dy.ptr(), x2=None, w=C.c_void_p(wT),
                             out=C.c_void_p(out.data_ptr() + 2 * n_off) if out_dtype else _p(out, n_off), bias=None,
                             emb=None, residual=residual, N=self.B, H=H, W=H, C1=128, C2=0, Ho=H, Wo=H, R=R, S=R, stride=1,
                             pad=pad, mode=0, w_cout=128, w_cin=128, n_off=0, out_ld=out_ld, emb_ld=0, res_ld=res_ld,
                             x_ld=128 if bf else dy.ld, x2_ld=0, round_out=1, op_dtype=1 if bf else 0, out_dtype=out_dtype)
            op = Op("b200_conv_igemm", a)
            op.tag = "dgrad"
            return op
        a = lib.ConvArgs(x=_p(dy.bf) if bf else dy.ptr(), x2=None, w=(self._wbf if bf else self._wtc)(wname + ".weight"),
                         out=C.c_void_p(out.data_ptr() + 2 * n_off) if out_dtype else _p(out, n_off), bias=None, emb=None,
                         residual=residual, N=self.B, H=H, W=H, C1=128, C2=0, Ho=H, Wo=H, R=R, S=R, stride=1, pad=pad,
                         mode=1, w_cout=128, w_cin=w_cin, n_off=n_off, out_ld=out_ld, emb_ld=0, res_ld=res_ld,
                         x_ld=128 if bf else dy.ld, x2_ld=0, round_out=1, op_dtype=1 if bf else 0, out_dtype=out_dtype)
        return Op("b200_conv_igemm", a)

    def _wgrad(self, x, x2, dy, wname, H, Ho, R, stride, pad, C1, C2, bf=False):
        a = lib.WgradArgs(x=_p(x), x2=_p(x2), dy=_p(dy.bf) if bf else dy.ptr(), dw=self._g(wname + ".weight"), N=self.B,
                          H=H, W=H, C1=C1, C2=C2, Ho=Ho, Wo=Ho, R=R, S=R, stride=stride, pad=pad, k_slices=0, x_ld=0,
                          x2_ld=0, dy_ld=128 if bf else dy.ld, op_dtype=1 if bf else 0)
        return Op("b200_conv_wgrad", a)

    def _gn(self, x, x2, gname, y, HW, C2, stats, silu=1, tf32=1, bf=False, x_bf=False, drop=False):
        a = lib.GnArgs(x=_p(x), x2=_p(x2), stats=_p(stats), gamma=self._w(gname + ".weight"),
                       beta=self._w(gname + ".bias"), y=_p(y), N=self.B, HW=HW, C1=128, C2=C2, x_ld=0, x2_ld=0,
                       eps=GN_EPS, silu=silu, round_tf32=tf32, y_dtype=1 if bf else 0, x_dtype=1 if x_bf else 0)
        if drop and self.dropout > 0.0:
            a.dropout_p = self.dropout
            a.dropout_seed = (zlib.crc32(gname.encode()) ^ self.drop_seed) & 0xFFFFFFFF
            a.dropout_step = _p(self.drop_step)
        return a

    def _ensure_bf(self, ops, g, HW):
        """bf16 mode: make sure the 128-channel gradient slice `g` has its dense bf16 copy (producers that are not a
        GroupNorm backward — upsample backward, plain-conv dgrad — get one cast pass)."""
        if g.bf is None:
            g.bf = self._get_bf(self.bfpool, self.B, HW, 128)
            ops.append(Op("b200_cast_bf16", lib.CastArgs(src=g.ptr(), dst=_p(g.bf), rows=self.B * HW, cols=128, src_ld=g.ld)))

    def _linear(self, mode, M, N, K, act, X=None, W=None, bias=None, Y=None, dY=None, dX=None, dW=None, db=None):
        a = lib.LinearArgs(X=_p(X), W=W, bias=bias, Y=_p(Y), dY=_p(dY), dX=_p(dX), dW=dW, db=db, M=M, N=N, K=K,
                           act=act, mode=mode)
        return Op("b200_linear", a)

    # ------------------------------------------------------------------ self-attention (extension; csrc/attn_small.cu)
    def _bgemm(self, A, Bm, Cm, M, N, K, lda, ldb, ldc, sA, sB, sC, a_mn=0, b_mn=0, alpha=1.0, round_out=0):
        a = lib.BgemmArgs(A=_p(A), B=_p(Bm), C=_p(Cm), M=M, N=N, K=K, batch=self.B, lda=lda, ldb=ldb, ldc=ldc, strideA=sA,
                          strideB=sB, strideC=sC, a_mn=a_mn, b_mn=b_mn, alpha=alpha, round_out=round_out)
        return Op("b200_bgemm", a)

    def _attn_fwd(self, ops, x, H):
        """x + proj(softmax(q k^T / sqrt(C)) v) over the H*H tokens of each sample, q / k / v / proj = 1x1 convolutions of
        GroupNorm(x) (oracle.attention_block).  Everything stays fp32 / TF32, also in bf16 mode (0.3 % of the FLOPs)."""
        import math
        B, T = self.B, H * H
        Tp = (T + 31) // 32 * 32
        pool = self.pool
        hn = pool.get(B, T, 128)
        g = self._gn(x, None, "attn.norm", hn, T, 0, self._next_stats(), silu=0, tf32=1)
        ops.append(Op("b200_gn_fwd", g))
        q, k, v = (pool.get(B, T, 128) for _ in range(3))
        for nm, t in (("q", q), ("k", k), ("v", v)):
            ops.append(self._conv(hn, None, "attn." + nm, t, H, H, 1, 1, 0, 128, 0))
        S, P = pool.get(B, T, Tp), pool.get(B, T, Tp)
        scale = 1.0 / math.sqrt(128.0)
        ops.append(self._bgemm(q, k, S, T, T, 128, 128, 128, Tp, T * 128, T * 128, T * Tp, alpha=scale))        # S = q k^T / sqrt(C)
        ops.append(Op("b200_softmax_fwd", lib.SoftmaxArgs(s=_p(S), p=_p(P), dp=None, ds=None, rows=B * T, T=T, ld=Tp)))
        a = pool.get(B, T, 128)
        ops.append(self._bgemm(P, v, a, T, 128, T, Tp, 128, 128, T * Tp, T * 128, T * 128, b_mn=1, round_out=1))  # a = P v
        out = pool.get(B, T, 128)
        out_bf = self._get_bf(pool, B, T, 128) if self.bf16 else None
        ops.append(self._conv(a, None, "attn.proj_out", out, H, H, 1, 1, 0, 128, 0, residual=_p(x), out_bf=out_bf))
        for t in (S, hn, q, k, v, P, a):
            pool.release(t)  # inference plans recycle; training plans keep everything
        self.records["attn"] = dict(x=x, H=H, Tp=Tp, hn=hn, q=q, k=k, v=v, P=P, a=a, g=g, scale=scale)
        return out

    def _attn_bwd(self, ops, dOut):
        """Reverse of _attn_fwd.  dOut: _Grad w.r.t. the block output; returns the dense gradient of its input."""
        r = self.records["attn"]
        B, H, Tp, scale = self.B, r["H"], r["Tp"], r["scale"]
        T = H * H
        gp = self.gpool
        q, k, v, P = r["q"], r["k"], r["v"], r["P"]
        # proj_out (1x1 conv): bias, weight and data gradients
        self._colsum(ops, dOut, T, total1=self._g("attn.proj_out.bias"))
        ops.append(self._wgrad(r["a"], None, dOut, "attn.proj_out", H, H, 1, 1, 0, 128, 0))
        dA = gp.get(B, T, 128)
        ops.append(self._dgrad(dOut, "attn.proj_out", dA, H, 1, 0, 128, 0, 128))
        # attention core: dV = P^T dA, dP = dA v^T, dS = softmax'(P, dP), dQ = dS k / sqrt(C), dK = dS^T q / sqrt(C)
        dV, dQ, dK = (gp.get(B, T, 128) for _ in range(3))
        dP, dS = gp.get(B, T, Tp), gp.get(B, T, Tp)
        ops.append(self._bgemm(P, dA, dV, T, 128, T, Tp, 128, 128, T * Tp, T * 128, T * 128, a_mn=1, b_mn=1, round_out=1))
        ops.append(self._bgemm(dA, v, dP, T, T, 128, 128, 128, Tp, T * 128, T * 128, T * Tp))
        ops.append(Op("b200_softmax_bwd", lib.SoftmaxArgs(s=None, p=_p(P), dp=_p(dP), ds=_p(dS), rows=B * T, T=T, ld=Tp)))
        ops.append(self._bgemm(dS, k, dQ, T, 128, T, Tp, 128, 128, T * Tp, T * 128, T * 128, b_mn=1, alpha=scale, round_out=1))
        ops.append(self._bgemm(dS, q, dK, T, 128, T, Tp, 128, 128, T * Tp, T * 128, T * 128, a_mn=1, b_mn=1, alpha=scale,
                               round_out=1))
        gp.release(dA), gp.release(dP), gp.release(dS)
        # q / k / v convs: their input gradients add up in dhn
        dhn = None
        for nm, gq in (("q", dQ), ("k", dK), ("v", dV)):
            gg = _Grad(gq)
            self._colsum(ops, gg, T, total1=self._g(f"attn.{nm}.bias"))
            ops.append(self._wgrad(r["hn"], None, gg, "attn." + nm, H, H, 1, 1, 0, 128, 0))
            nxt = gp.get(B, T, 128)
            ops.append(self._dgrad(gg, "attn." + nm, nxt, H, 1, 0, 128, 0, 128,
                                   residual=_p(dhn) if dhn is not None else None, res_ld=128))
            gp.release(dhn), gp.release(gq)
            dhn = nxt
        dx = gp.get(B, T, 128)
        self._gn_bwd(ops, r["g"], dhn, dx, "attn.norm", add1=dOut, add_c=128)  # + the residual path's gradient
        gp.release(dhn)
        return dx

    # ------------------------------------------------------------------ forward
    def _build_embed(self, force_uncond):
        """simple_unet.py:45-58 + the 12 emb_layers (:165-167,183) as one [B,256]x[256,1536] GEMM."""
        B, E = self.B, self.E
        st = self.store
        ops = []
        ea = lib.EmbedArgs(t=_p(self.t) if self.sched is None else None, sched=_p(self.sched),
                           step_dev=_p(self.step_dev), freqs=_p(self.freqs), guide=_p(self.guide),
                           sincos=_p(self.sincos), onehot=_p(self.onehot), t_out=_p(self.t) if self.sched is not None else None,
                           t_emb=_p(self.temb), g_emb=_p(self.gemb), emb=_p(self.emb), semb=_p(self.semb),
                           d_semb=None, d_emb=None, d_g_emb=None, B=B, E=E, step=0, force_uncond=force_uncond)
        ops.append(Op("b200_embed_prep", ea))
        ops.append(self._linear(0, B, E, 64, 0, X=self.sincos, W=self._w("time_embed.0.weight"),
                                bias=self._w("time_embed.0.bias"), Y=self.te0))
        ops.append(self._linear(0, B, E, E, 1, X=self.te0, W=self._w("time_embed.2.weight"),
                                bias=self._w("time_embed.2.bias"), Y=self.temb))
        if not force_uncond:
            ops.append(self._linear(0, B, E, 10, 0, X=self.onehot, W=self._w("guide_embed.0.weight"),
                                    bias=self._w("guide_embed.0.bias"), Y=self.ge0))
            ops.append(self._linear(0, B, E, E, 1, X=self.ge0, W=self._w("guide_embed.2.weight"),
                                    bias=self._w("guide_embed.2.bias"), Y=self.gemb))
        if self.use_cond_w:
            # simple_unet.py:60-64: emb += cond_w_embed(timestep_embedding(cond_w, 64, max_period=4))
            wa = lib.EmbedArgs(t=_p(self.cond_w), sched=None, step_dev=None, freqs=_p(self.freqs_w), guide=None,
                               sincos=_p(self.sincos_w), onehot=None, t_out=None, t_emb=None, g_emb=None, emb=None, semb=None,
                               d_semb=None, d_emb=None, d_g_emb=None, B=B, E=E, step=0, force_uncond=0)
            ops.append(Op("b200_embed_prep", wa))
            ops.append(self._linear(0, B, E, 64, 0, X=self.sincos_w, W=self._w("cond_w_embed.0.weight"),
                                    bias=self._w("cond_w_embed.0.bias"), Y=self.cw0))
            ops.append(self._linear(0, B, E, E, 1, X=self.cw0, W=self._w("cond_w_embed.2.weight"),
                                    bias=self._w("cond_w_embed.2.bias"), Y=self.cwemb))
            ea.w_emb = _p(self.cwemb)
        ops.append(Op("b200_embed_combine", ea))
        ops.append(self._linear(0, B, 12 * self.Cc, E, 0, X=self.semb, W=C.c_void_p(st.flat.data_ptr() + 4 * st.emb_w_off),
                                bias=C.c_void_p(st.flat.data_ptr() + 4 * st.emb_b_off), Y=self.embout))
        return ops

    def _rb_fwd(self, ops, name, x, x2, H):
        """ResBlock (simple_unet.py:155-186): GN-SiLU-conv3x3 (+emb) - GN-SiLU-conv3x3 (+skip)."""
        B, HW = self.B, H * H
        k = RESBLOCKS.index(name)
        C2 = 128 if x2 is not None else 0
        Cin = 128 + C2
        pool = self.pool
        st1, st2 = self._next_stats(), self._next_stats()
        bf = self.bf16
        a1 = self._get_bf(pool, B, HW, Cin) if bf else pool.get(B, HW, Cin)
        g1 = self._gn(x, x2, name + ".in_layers.0", a1, HW, C2, st1, bf=bf)
        ops.append(Op("b200_gn_fwd", g1))
        # h1 feeds only the GroupNorm (never a tensor-core operand): full fp32 mantissa in tf32 mode, bf16 storage in
        # bf16 mode (like the reference's autocast conv output)
        bfs = self.bfs
        h1 = self._get_bf(pool, B, HW, 128) if bfs else pool.get(B, HW, 128)
        ops.append(self._conv(a1, None, name + ".in_layers.2", h1, H, H, 3, 1, 1, Cin, 0,
                              emb=_p(self.embout, k * 128), round_out=0, bf=bf, out_dtype=1 if bfs else 0))
        pool.release(a1)
        a2 = self._get_bf(pool, B, HW, 128) if bf else pool.get(B, HW, 128)
        # (the reference builds up.seq.3's ResBlock without its dropout argument, simple_unet.py:137-139: p = 0 there)
        g2 = self._gn(h1, None, name + ".out_layers.0", a2, HW, 0, st2, bf=bf, x_bf=bfs, drop=(name != "up.seq.3.0"))
        ops.append(Op("b200_gn_fwd", g2))
        pool.release(h1)
        skip = None
        x_bf = x2_bf = None
        if x2 is not None:
            skip = pool.get(B, HW, 128)
            if bf:  # the 1x1 skip conv reads the bf16 copies of the running tensor and of the cached Down tensor
                x_bf, x2_bf = x.bf16_copy, x2.bf16_copy
                ops.append(self._conv(x_bf, x2_bf, name + ".skip_connection", skip, H, H, 1, 1, 0, 128, 128, bf=True))
            else:
                ops.append(self._conv(x, x2, name + ".skip_connection", skip, H, H, 1, 1, 0, 128, 128))
            residual = skip
        else:
            residual = x
        out = pool.get(B, HW, 128)
        out_bf = self._get_bf(pool, B, HW, 128) if bf else None
        ops.append(self._conv(a2, None, name + ".out_layers.3", out, H, H, 3, 1, 1, 128, 0, residual=_p(residual), bf=bf,
                              out_bf=out_bf))
        pool.release(a2)
        pool.release(skip)
        self.records[name] = dict(x=x, x2=x2, H=H, a1=a1, h1=h1, a2=a2, g1=g1, g2=g2, k=k, x_bf=x_bf, x2_bf=x2_bf)
        return out

    def _build_forward(self):
        B, S, E = self.B, self.S, self.E
        dev = self.store.device
        f32 = dict(dtype=torch.float32, device=dev)
        for nm, w in (("sincos", 64), ("onehot", 10), ("te0", E), ("temb", E), ("ge0", E), ("gemb", E), ("emb", E),
                      ("semb", E), ("embout", 12 * self.Cc), ("sincos_w", 64), ("cw0", E), ("cwemb", E)):
            setattr(self, nm, torch.zeros(B, w, **f32))
        self.embed_ops = [self._build_embed(0), self._build_embed(1)]
        st = self.store
        self.refresh_ops = [Op("b200_round_tf32", lib.RoundArgs(src=_p(st.flat), dst=_p(st.flat_tc), n=st.numel))]
        if self.bf16:
            self.refresh_ops.append(Op("b200_cast_bf16", lib.CastArgs(src=_p(st.flat), dst=_p(st.flat_bf), rows=1,
                                                                      cols=st.numel, src_ld=0)))
        ops = []
        pool = self.pool
        H0, H1, H2 = S, S // 2, S // 4
        # ---- Down (simple_unet.py:87-109)
        d0 = pool.get(B, H0 * H0, 128)
        Ci = self.Ci
        for ci in range(Ci):  # one launch per image channel; launches after the first accumulate onto d0
            cin = lib.ConvInArgs(x=_p(self.z_in, ci * H0 * H0), w=C.c_void_p(st.ptr("down.seq.0.conv.weight") + 4 * ci),
                                 bias=self._w("down.seq.0.conv.bias"), out=_p(d0), dy=None, dw=None, db=None, N=B, H=H0,
                                 W=H0, round_out=1 if ci == Ci - 1 else 0, x_ld=Ci * H0 * H0, w_co_ld=9 * Ci, w_tap_ld=Ci,
                                 accumulate=1 if ci else 0)
            ops.append(Op("b200_conv_in_fwd", cin))
        bf = self.bf16
        if bf:
            self._cast_copy(ops, d0, B * H0 * H0)
        d1 = self._rb_fwd(ops, "down.seq.1", d0, None, H0)
        d2 = self._rb_fwd(ops, "down.seq.2", d1, None, H0)
        d3 = pool.get(B, H1 * H1, 128)
        ops.append(self._conv(d2, None, "down.seq.3.conv", d3, H0, H1, 3, 2, 1, 128, 0,
                              out_bf=self._get_bf(pool, B, H1 * H1, 128) if bf else None))
        d4 = self._rb_fwd(ops, "down.seq.4", d3, None, H1)
        d5 = self._rb_fwd(ops, "down.seq.5", d4, None, H1)
        d6 = pool.get(B, H2 * H2, 128)
        ops.append(self._conv(d5, None, "down.seq.6.conv", d6, H1, H2, 3, 2, 1, 128, 0,
                              out_bf=self._get_bf(pool, B, H2 * H2, 128) if bf else None))
        # ---- turn (simple_unet.py:36,68)
        t = self._rb_fwd(ops, "turn", d6, None, H2)
        if self.attention:
            t_in = t
            t = self._attn_fwd(ops, t_in, H2)
            self._rel(t_in)
        # ---- Up (simple_unet.py:125-152): block i consumes cat([h, cache.pop()])
        r0 = self._rb_fwd(ops, "up.seq.0.0", t, d6, H2)
        self._rel(t), self._rel(d6)
        # nearest 2x upsample feeds only the following conv (fprop + wgrad): in bf16 mode it is written as bf16
        up0 = self._get_bf(pool, B, H1 * H1, 128) if bf else pool.get(B, H1 * H1, 128)
        ops.append(Op("b200_upsample2x", lib.ResampleArgs(in_=_p(r0), out=_p(up0), N=B, H=H2, W=H2, C=128, in_ld=0, mode=0,
                                                          out_dtype=1 if bf else 0)))
        self._rel(r0)
        u0 = pool.get(B, H1 * H1, 128)
        ops.append(self._conv(up0, None, "up.seq.0.1.conv", u0, H1, H1, 3, 1, 1, 128, 0, bf=bf,
                              out_bf=self._get_bf(pool, B, H1 * H1, 128) if bf else None))
        pool.release(up0)
        u1 = self._rb_fwd(ops, "up.seq.1", u0, d5, H1)
        self._rel(u0), self._rel(d5)
        u2 = self._rb_fwd(ops, "up.seq.2", u1, d4, H1)
        self._rel(u1), self._rel(d4)
        r3 = self._rb_fwd(ops, "up.seq.3.0", u2, d3, H1)
        self._rel(u2), self._rel(d3)
        # nearest 2x upsample feeds only the following conv (fprop + wgrad): in bf16 mode it is written as bf16
        up3 = self._get_bf(pool, B, H0 * H0, 128) if bf else pool.get(B, H0 * H0, 128)
        ops.append(Op("b200_upsample2x", lib.ResampleArgs(in_=_p(r3), out=_p(up3), N=B, H=H1, W=H1, C=128, in_ld=0, mode=0,
                                                          out_dtype=1 if bf else 0)))
        self._rel(r3)
        u3 = pool.get(B, H0 * H0, 128)
        ops.append(self._conv(up3, None, "up.seq.3.1.conv", u3, H0, H0, 3, 1, 1, 128, 0, bf=bf,
                              out_bf=self._get_bf(pool, B, H0 * H0, 128) if bf else None))
        pool.release(up3)
        u4 = self._rb_fwd(ops, "up.seq.4", u3, d2, H0)
        self._rel(u3), self._rel(d2)
        u5 = self._rb_fwd(ops, "up.seq.5", u4, d1, H0)
        self._rel(u4), self._rel(d1)
        u6 = self._rb_fwd(ops, "up.seq.6", u5, d0, H0)
        self._rel(u5), self._rel(d0)
        # ---- out head (simple_unet.py:38-42,70)
        sth = self._next_stats()
        ah = pool.get(B, H0 * H0, 128)
        gh = self._gn(u6, None, "out.0", ah, H0 * H0, 0, sth, silu=1, tf32=0)
        ops.append(Op("b200_gn_fwd", gh))
        self.body_ops = ops
        self.head_ops = []
        for o in (self.out, self.out2):  # head_ops[k] = the launches (one per image channel) that write output buffer k
            grp = []
            for ci in range(Ci):
                ca = lib.ConvOutArgs(h=_p(ah), w=C.c_void_p(st.ptr("out.2.weight") + 4 * ci * 1152),
                                     bias=C.c_void_p(st.ptr("out.2.bias") + 4 * ci), out=_p(o, ci * H0 * H0), dout=None,
                                     dh=None, dw=None, db=None, N=B, H=H0, W=H0, out_ld=Ci * H0 * H0)
                grp.append(Op("b200_conv_out_fwd", ca))
            self.head_ops.append(grp)
        self.saved = dict(d0=d0, d1=d1, d2=d2, d3=d3, d4=d4, d5=d5, d6=d6, t=t, r0=r0, up0=up0, u0=u0, u1=u1, u2=u2,
                          r3=r3, up3=up3, u3=u3, u4=u4, u5=u5, u6=u6, ah=ah, gh=gh)

    def forward(self, stream_ptr, uncond=False, to_out2=False, refresh=True):
        """Enqueue embeddings + UNet body + head.  Inputs are the static buffers z_in / t (or sched+step_dev) / guide.
        refresh: re-derive the TF32 weight shadow first (one 24 MB pass; skip it when the weights cannot have changed)."""
        if refresh:
            run_ops(self.refresh_ops, stream_ptr)
        run_ops(self.embed_ops[1 if uncond else 0], stream_ptr)
        run_ops(self.body_ops, stream_ptr)
        run_ops(self.head_ops[1 if to_out2 else 0], stream_ptr)

    # ------------------------------------------------------------------ backward
    def _gn_bwd(self, ops, g_fwd, dy, dx, gname, add1=None, add2=None, add_c=0, dx_bf=None, dy_bf=False):
        """dx_bf: storage for the bf16 copy of dx[:, :128] (bf16 mode: the dY operand of the conv that produced x)."""
        a = lib.GnBwdArgs(fwd=g_fwd, dy=_p(dy), dx=_p(dx), add1=add1.ptr() if add1 else None,
                          add2=add2.ptr() if add2 else None, add1_ld=add1.ld if add1 else 0,
                          add2_ld=add2.ld if add2 else 0, add_c=add_c, dgamma=self._g(gname + ".weight"),
                          dbeta=self._g(gname + ".bias"), round_dx=1, dx_bf16=_p(dx_bf),
                          dx_bf16_c=128 if dx_bf is not None else 0, dy_dtype=1 if dy_bf else 0)
        ops.append(Op("b200_gn_bwd", a, ws=self.gn_ws))
        if dx is None:  # bf16-only output
            dx_bf.producer_gn = a
            return
        dx.producer_gn = a
        if dx_bf is not None:
            dx.bf16_copy = dx_bf

    def _colsum(self, ops, dy, HW, per_sample=None, ps_ld=0, total1=None, total2=None):
        g = dy.prod
        if g is not None and not g.colsum_c and self.fuse_colsum:
            g.colsum_per_sample, g.colsum_ps_ld, g.colsum_total1, g.colsum_total2 = per_sample, ps_ld, total1, total2
            g.colsum_c = 128
            return
        assert dy.t is not dy.bf, "a bf16-only gradient has no fp32 tensor for the stand-alone colsum kernel"
        a = lib.ColsumArgs(dy=dy.ptr(), per_sample=per_sample, total1=total1, total2=total2, N=self.B, HW=HW, ld=dy.ld,
                           ps_ld=ps_ld)
        ops.append(Op("b200_colsum", a))

    def _rb_bwd(self, ops, name, dOut, extra=None):
        """Reverse of _rb_fwd.  dOut: _Grad w.r.t. the block output; extra: _Grad that also flows into the block
        input (the skip-concat consumer of a Down tensor).  Returns the dense [B,HW,Cin] gradient of the input."""
        r = self.records[name]
        B, H = self.B, r["H"]
        HW = H * H
        x, x2 = r["x"], r["x2"]
        Cin = 256 if x2 is not None else 128
        gp = self.gpool
        has_skip = x2 is not None
        bf = self.bf16
        if bf:
            self._ensure_bf(ops, dOut, HW)
        # conv2 (out_layers.3): bias, weight, data gradients
        self._colsum(ops, dOut, HW, total1=self._g(name + ".out_layers.3.bias"),
                     total2=self._g(name + ".skip_connection.bias") if has_skip else None)
        ops.append(self._wgrad(r["a2"], None, dOut, name + ".out_layers.3", H, H, 3, 1, 1, 128, 0, bf=bf))
        # bf16 mode: the dgrad outputs (read only by the GroupNorm backward) and dh1 (read only by conv1's dgrad /
        # wgrad; its column sums are fused into the GroupNorm backward) are stored as bf16 only
        bfs = self.bfs
        da2 = self._get_bf(gp, B, HW, 128) if bfs else gp.get(B, HW, 128)
        ops.append(self._dgrad(dOut, name + ".out_layers.3", da2, H, 3, 1, 128, 0, 128, bf=bf, out_dtype=1 if bfs else 0))
        dh1 = None if self.dh1_bf_only else gp.get(B, HW, 128)
        dh1_bf = self._get_bf(gp, B, HW, 128) if bf else None
        self._gn_bwd(ops, r["g2"], da2, dh1, name + ".out_layers.0", dx_bf=dh1_bf, dy_bf=bfs)
        gp.release(da2)
        # emb_out broadcast add (simple_unet.py:183-184) and conv1 bias
        dh1g = _Grad(dh1_bf, 0, 128, bf=dh1_bf) if dh1 is None else _Grad(dh1)
        self._colsum(ops, dh1g, HW, per_sample=_p(self.d_embout, r["k"] * 128), ps_ld=12 * self.Cc,
                     total1=self._g(name + ".in_layers.2.bias"))
        ops.append(self._wgrad(r["a1"], None, dh1g, name + ".in_layers.2", H, H, 3, 1, 1, Cin, 0, bf=bf))
        da1 = self._get_bf(gp, B, HW, Cin) if bfs else gp.get(B, HW, Cin)
        for n_off in range(0, Cin, 128):
            ops.append(self._dgrad(dh1g, name + ".in_layers.2", da1, H, 3, 1, Cin, n_off, Cin, bf=bf,
                                   out_dtype=1 if bfs else 0))
        gp.release(dh1)
        gp.release(dh1_bf)
        dsk = None
        if has_skip:
            dsk = gp.get(B, HW, 256)
            for n_off in (0, 128):
                ops.append(self._dgrad(dOut, name + ".skip_connection", dsk, H, 1, 0, 256, n_off, 256, bf=bf))
            if bf:
                ops.append(self._wgrad(r["x_bf"], r["x2_bf"], dOut, name + ".skip_connection", H, H, 1, 1, 0, 128, 128, bf=True))
            else:
                ops.append(self._wgrad(x, x2, dOut, name + ".skip_connection", H, H, 1, 1, 0, 128, 128))
            add1, add_c = _Grad(dsk), 256
        else:
            add1, add_c = dOut, 128
        dx = (self.pool_dcat if has_skip else gp).get(B, HW, Cin)
        # the first 128 channels of dx are the output gradient of whatever produced x: give it its bf16 copy here
        dx_bf = self._get_bf(self.bfpool, B, HW, 128) if bf else None
        self._gn_bwd(ops, r["g1"], da1, dx, name + ".in_layers.0", add1=add1, add2=extra, add_c=add_c, dx_bf=dx_bf,
                     dy_bf=bfs)
        gp.release(da1)
        gp.release(dsk)
        return dx

    def _conv_bwd(self, ops, name, x, dy, H, Ho, stride, residual=None):
        """Backward of a plain 3x3 conv layer (Downsample / Upsample conv).  dy: _Grad at the output resolution.
        Returns the dense input gradient (+ residual: another _Grad flowing into the same tensor)."""
        B = self.B
        gp = self.gpool
        bf = self.bf16 and stride == 1  # the Upsample convs (x is the bf16 upsampled tensor); Downsample convs stay TF32
        if bf:
            self._ensure_bf(ops, dy, Ho * Ho)
        self._colsum(ops, dy, Ho * Ho, total1=self._g(name + ".bias"))
        ops.append(self._wgrad(x, None, dy, name, H, Ho, 3, stride, 1, 128, 0, bf=bf))
        src, zs = dy, None
        if stride == 2 and self.dgrad_T and self.dgrad_s2:
            # mode 2: the four output-parity classes of dX as small forward convs over dY (9 taps instead of 36)
            dx = gp.get(B, H * H, 128)
            a = lib.ConvArgs(x=dy.ptr(), x2=None, w=C.c_void_p(self.store.tcT_ptr(name + ".weight")), out=_p(dx), bias=None,
                             emb=None, residual=residual.ptr() if residual else None, N=B, H=Ho, W=Ho, C1=128, C2=0, Ho=H,
                             Wo=H, R=3, S=3, stride=2, pad=1, mode=2, w_cout=128, w_cin=128, n_off=0, out_ld=128,
                             emb_ld=0, res_ld=residual.ld if residual else 0, x_ld=dy.ld, x2_ld=0, round_out=1)
            op = Op("b200_conv_igemm", a)
            op.tag = "dgrad"
            op.alg_scale = 0.25  # conv_flops counts M = B*H*H output pixels; each sees 9/4 taps on average
            ops.append(op)
            return dx
        if stride == 2:
            # zero-stuff dY to the input resolution so the stride-1 dgrad kernel applies
            zs = gp.get(B, H * H, 128)
            ops.append(Op("b200_upsample2x", lib.ResampleArgs(in_=dy.ptr(), out=_p(zs), N=B, H=H, W=H, C=128,
                                                              in_ld=dy.ld, mode=2)))
            src = _Grad(zs)
        dx = gp.get(B, H * H, 128)
        ops.append(self._dgrad(src, name, dx, H, 3, 1, 128, 0, 128,
                               residual=residual.ptr() if residual else None, res_ld=residual.ld if residual else 0, bf=bf))
        if stride == 2:
            ops[-1].alg_scale = 0.25  # 3/4 of the zero-stuffed rows are zeros
            if self.bf16 and self.dgrad_T:  # the ResBlock below reads dx as a bf16 dY operand: no cast pass
                dx.bf16_copy = self._get_bf(self.bfpool, B, H * H, 128)
                ops[-1].args.out_bf16 = _p(dx.bf16_copy)
        gp.release(zs)
        return dx

    def _build_backward(self):
        B, S, E = self.B, self.S, self.E
        st = self.store
        dev = st.device
        f32 = dict(dtype=torch.float32, device=dev)
        H0, H1, H2 = S, S // 2, S // 4
        sv = self.saved
        gp = self.gpool
        import os
        self.dgrad_T = os.environ.get("B200_DGRAD_T", "1") == "1"
        # Downsample dgrad: 0 (default) = zero-stuff dY + the stride-1 strip kernel; 1 = conv_igemm mode 2 (parity
        # classes, no zero-stuffing: 9 taps instead of 36, but one small epilogue-dominated CTA per class and tile —
        # measured 301 vs 331 us at 28x28 and 92 vs 50 us at 14x14 with the residual add, so not the default)
        self.dgrad_s2 = os.environ.get("B200_DGRAD_S2", "0") == "1"
        self.refreshT_ops = []
        if self.dgrad_T:
            table = st.ensure_transposed(self.bf16)
            self.refreshT_ops.append(Op("b200_conv_weights_transpose", lib.WTransposeArgs(
                src=_p(st.flat), dst=_p(st.flat_tcT), table=_p(table), n_weights=table.shape[0], max_elems=st.wT_max,
                out_dtype=0)))
            if self.bf16:
                self.refreshT_ops.append(Op("b200_conv_weights_transpose", lib.WTransposeArgs(
                    src=_p(st.flat), dst=_p(st.flat_bfT), table=_p(table), n_weights=table.shape[0], max_elems=st.wT_max,
                    out_dtype=1)))
        self.pool_dcat = _Pool(dev, recycle=False)  # concat gradients live until the Down path consumes their 2nd half
        self.bfpool = _Pool(dev, recycle=False)     # bf16 copies of block-output gradients (live across blocks)
        # one workspace shared by every b200_gn_bwd launch of the plan, sized by the library's own query
        probe = lib.GnBwdArgs(fwd=lib.GnArgs(N=B, HW=S * S, C1=128, C2=128))
        self.gn_ws = torch.zeros(max(1, lib.load().b200_gn_bwd_workspace(C.byref(probe)) // 4), **f32)
        self.d_embout = torch.zeros(B, 12 * self.Cc, **f32)
        for nm in ("d_semb", "d_emb", "d_gemb", "d_te0", "d_ge0"):
            setattr(self, nm, torch.zeros(B, E, **f32))
        ops = []
        ops.append(Op("b200_zero", lib.ZeroArgs(ptr=_p(self.d_embout), bytes=self.d_embout.numel() * 4)))
        # ---- out head
        dah = gp.get(B, H0 * H0, 128)
        Ci = self.Ci
        for ci in range(Ci):
            ca = lib.ConvOutArgs(h=_p(sv["ah"]), w=C.c_void_p(st.ptr("out.2.weight") + 4 * ci * 1152),
                                 bias=self._w("out.2.bias"), out=None, dout=_p(self.dout, ci * H0 * H0), dh=_p(dah),
                                 dw=C.c_void_p(st.gptr("out.2.weight") + 4 * ci * 1152),
                                 db=C.c_void_p(st.gptr("out.2.bias") + 4 * ci), N=B, H=H0, W=H0,
                                 round_dh=1 if ci == Ci - 1 else 0, out_ld=Ci * H0 * H0, accumulate_dh=1 if ci else 0)
            ops.append(Op("b200_conv_out_bwd", ca))
        du6 = gp.get(B, H0 * H0, 128)
        self._gn_bwd(ops, sv["gh"], dah, du6, "out.0",
                     dx_bf=self._get_bf(self.bfpool, B, H0 * H0, 128) if self.bf16 else None)
        gp.release(dah)
        # ---- Up, reversed
        half = lambda t: _Grad(t, 0, 256)       # gradient of the running tensor h
        tail = lambda t: _Grad(t, 128, 256)     # gradient of the cached Down tensor
        dc6 = self._rb_bwd(ops, "up.seq.6", _Grad(du6))
        gp.release(du6)
        dc5 = self._rb_bwd(ops, "up.seq.5", half(dc6))
        dc4 = self._rb_bwd(ops, "up.seq.4", half(dc5))
        dup3 = self._conv_bwd(ops, "up.seq.3.1.conv", sv["up3"], half(dc4), H0, H0, 1)
        dr3 = gp.get(B, H1 * H1, 128)
        ops.append(Op("b200_upsample2x", lib.ResampleArgs(in_=_p(dup3), out=_p(dr3), N=B, H=H1, W=H1, C=128, in_ld=0, mode=1, round_out=1)))
        gp.release(dup3)
        dc3 = self._rb_bwd(ops, "up.seq.3.0", _Grad(dr3))
        gp.release(dr3)
        dc2 = self._rb_bwd(ops, "up.seq.2", half(dc3))
        dc1 = self._rb_bwd(ops, "up.seq.1", half(dc2))
        dup0 = self._conv_bwd(ops, "up.seq.0.1.conv", sv["up0"], half(dc1), H1, H1, 1)
        dr0 = gp.get(B, H2 * H2, 128)
        ops.append(Op("b200_upsample2x", lib.ResampleArgs(in_=_p(dup0), out=_p(dr0), N=B, H=H2, W=H2, C=128, in_ld=0, mode=1, round_out=1)))
        gp.release(dup0)
        dc0 = self._rb_bwd(ops, "up.seq.0.0", _Grad(dr0))
        gp.release(dr0)
        # every gradient of out.* and up.* is final here (ParamStore.bucket_split): data-parallel training starts the
        # allreduce of that bucket while the ops below (turn, Down, embeddings) still run
        self.bwd_split = len(ops)
        # ---- turn, then Down reversed; each Down tensor also receives the tail of its concat consumer
        if self.attention:
            dt = self._attn_bwd(ops, half(dc0))
            dd6 = self._rb_bwd(ops, "turn", _Grad(dt), extra=tail(dc0))
            gp.release(dt)
        else:
            dd6 = self._rb_bwd(ops, "turn", half(dc0), extra=tail(dc0))
        dd5 = self._conv_bwd(ops, "down.seq.6.conv", sv["d5"], _Grad(dd6), H1, H2, 2, residual=tail(dc1))
        gp.release(dd6)
        dd4 = self._rb_bwd(ops, "down.seq.5", _Grad(dd5), extra=tail(dc2))
        gp.release(dd5)
        dd3 = self._rb_bwd(ops, "down.seq.4", _Grad(dd4), extra=tail(dc3))
        gp.release(dd4)
        dd2 = self._conv_bwd(ops, "down.seq.3.conv", sv["d2"], _Grad(dd3), H0, H1, 2, residual=tail(dc4))
        gp.release(dd3)
        dd1 = self._rb_bwd(ops, "down.seq.2", _Grad(dd2), extra=tail(dc5))
        gp.release(dd2)
        dd0 = self._rb_bwd(ops, "down.seq.1", _Grad(dd1), extra=tail(dc6))
        gp.release(dd1)
        for ci in range(Ci):
            cin = lib.ConvInArgs(x=_p(self.z_in, ci * H0 * H0), w=self._w("down.seq.0.conv.weight"), bias=None, out=None,
                                 dy=_p(dd0), dw=C.c_void_p(st.gptr("down.seq.0.conv.weight") + 4 * ci),
                                 db=self._g("down.seq.0.conv.bias") if ci == 0 else None, N=B, H=H0, W=H0,
                                 x_ld=Ci * H0 * H0, w_co_ld=9 * Ci, w_tap_ld=Ci)
            ops.append(Op("b200_conv_in_bwd", cin))
        gp.release(dd0)
        # ---- embeddings (simple_unet.py:45-58 and the emb_layers), reversed
        embw = C.c_void_p(st.flat.data_ptr() + 4 * st.emb_w_off)
        ops.append(self._linear(2, B, 12 * self.Cc, E, 0, X=self.semb, W=embw, dY=self.d_embout,
                                dW=C.c_void_p(st.grad.data_ptr() + 4 * st.emb_w_off),
                                db=C.c_void_p(st.grad.data_ptr() + 4 * st.emb_b_off)))
        ops.append(self._linear(1, B, 12 * self.Cc, E, 0, W=embw, dY=self.d_embout, dX=self.d_semb))
        ea = lib.EmbedArgs(t=None, sched=None, step_dev=None, freqs=_p(self.freqs), guide=_p(self.guide), sincos=None,
                           onehot=None, t_out=None, t_emb=None, g_emb=None, emb=_p(self.emb), semb=None,
                           d_semb=_p(self.d_semb), d_emb=_p(self.d_emb), d_g_emb=_p(self.d_gemb), B=B, E=E, step=0,
                           force_uncond=0)
        ops.append(Op("b200_embed_combine_bwd", ea))
        branches = [("time_embed", self.d_emb, self.sincos, 64, self.d_te0, self.te0),
                    ("guide_embed", self.d_gemb, self.onehot, 10, self.d_ge0, self.ge0)]
        if self.use_cond_w:  # d(emb)/d(cond_w embedding) = 1: its output gradient is d_emb itself
            self.d_cw0 = torch.zeros(B, E, **f32)
            branches.append(("cond_w_embed", self.d_emb, self.sincos_w, 64, self.d_cw0, self.cw0))
        for pre, dY, X0, K0, dmid, mid in branches:
            ops.append(self._linear(2, B, E, E, 1, X=mid, W=self._w(pre + ".2.weight"), dY=dY,
                                    dW=self._g(pre + ".2.weight"), db=self._g(pre + ".2.bias")))
            ops.append(self._linear(1, B, E, E, 1, X=mid, W=self._w(pre + ".2.weight"), dY=dY, dX=dmid))
            ops.append(self._linear(2, B, E, K0, 0, X=X0, W=self._w(pre + ".0.weight"), dY=dmid,
                                    dW=self._g(pre + ".0.weight"), db=self._g(pre + ".0.bias")))
        self.bwd_ops = ops

    def backward(self, stream_ptr, part=None):
        """Enqueue the reverse pass: reads self.dout, accumulates into store.grad (caller zeroes it once per step).
        part: None = all of it; "up" = out head + Up blocks (ops [0, bwd_split)); "down" = the rest."""
        if part in (None, "up"):
            run_ops(self.refreshT_ops, stream_ptr)  # the weights have not changed since the forward's refresh
        lo = self.bwd_split if part == "down" else 0
        hi = self.bwd_split if part == "up" else len(self.bwd_ops)
        run_ops(self.bwd_ops[lo:hi], stream_ptr)

    def activation_bytes(self):
        n = self.pool.bytes() + self.gpool.bytes()
        if self.train:
            n += self.pool_dcat.bytes() + self.bfpool.bytes()
        return n
