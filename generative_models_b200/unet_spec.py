"""Static description of the reference UNet (gms/diffusion/simple_unet.py:16-186): parameter names/shapes in the
reference's registration order (== its state_dict order), the 12 ResBlocks in execution order, and the flat fp32
parameter store whose views are the nn.Parameters.

Layout decisions (B200-first):
  * one flat fp32 buffer for all parameters (and one for gradients, two for Adam state): a single NCCL allreduce
    and a single fused Adam kernel cover the model;
  * conv weights are logical OIHW parameters with channels_last strides, i.e. physically [Cout][kh][kw][Cin] — the
    K-major B operand of the implicit GEMM — so state_dict() stays reference-compatible (Appendix C of SURVEY.md);
  * the 12 emb_layers Linear(256 -> 128) weights are contiguous, so they run as ONE [B,256]x[256,1536] GEMM.
"""
import math

import torch
from torch import nn

# ResBlocks in execution order; the index is the column block of the fused emb_layers GEMM.
RESBLOCKS = ["down.seq.1", "down.seq.2", "down.seq.4", "down.seq.5", "turn", "up.seq.0.0", "up.seq.1", "up.seq.2",
             "up.seq.3.0", "up.seq.4", "up.seq.5", "up.seq.6"]


def param_specs(C=128, img_channels=1, attention=False):
    """[(name, shape, kind)] in the reference's registration order; kind in {lin_w, lin_b, conv_w, conv_b, gn_w, gn_b}.
    `zero` convs (simple_unet.py:172 zero_module) are flagged in ZERO_INIT.  img_channels = 1 is the reference
    (simple_unet.py:92-94 Conv2d(1, C, 3), :41 Conv2d(C, 1, 3)); 3 is the 3x64x64 extension of BASELINE.json configs[4]."""
    E = 2 * C
    out = []

    def lin(name, i, o):
        out.append((name + ".weight", (o, i), "lin_w"))
        out.append((name + ".bias", (o,), "lin_b"))

    def conv(name, i, o, k):
        out.append((name + ".weight", (o, i, k, k), "conv_w"))
        out.append((name + ".bias", (o,), "conv_b"))

    def gn(name, c):
        out.append((name + ".weight", (c,), "gn_w"))
        out.append((name + ".bias", (c,), "gn_b"))

    def rb(pre, cin, cout):
        gn(pre + ".in_layers.0", cin)
        conv(pre + ".in_layers.2", cin, cout, 3)
        lin(pre + ".emb_layers.1", E, cout)
        gn(pre + ".out_layers.0", cout)
        conv(pre + ".out_layers.3", cout, cout, 3)
        if cin != cout:
            conv(pre + ".skip_connection", cin, cout, 1)

    for nm, i in (("time_embed", 64), ("cond_w_embed", 64), ("guide_embed", 10)):
        lin(nm + ".0", i, E)
        lin(nm + ".2", E, E)
    conv("down.seq.0.conv", img_channels, C, 3)
    rb("down.seq.1", C, C)
    rb("down.seq.2", C, C)
    conv("down.seq.3.conv", C, C, 3)
    rb("down.seq.4", C, C)
    rb("down.seq.5", C, C)
    conv("down.seq.6.conv", C, C, 3)
    rb("turn", C, C)
    if attention:  # extension (no reference counterpart): self-attention after the turn block, at the lowest resolution
        gn("attn.norm", C)
        for n in ("q", "k", "v", "proj_out"):
            conv("attn." + n, C, C, 1)
    for i in range(7):
        if i in (0, 3):
            rb(f"up.seq.{i}.0", 2 * C, C)
            conv(f"up.seq.{i}.1.conv", C, C, 3)
        else:
            rb(f"up.seq.{i}", 2 * C, C)
    gn("out.0", C)
    conv("out.2", C, img_channels, 3)
    return out


def is_zero_init(name):
    return ".out_layers.3." in name or name.startswith("attn.proj_out.")


def _align(n, a=64):
    return (n + a - 1) // a * a


class ParamStore:
    """Flat fp32 parameter / gradient / Adam-state buffers on one device, plus named views."""

    def __init__(self, C, device, img_channels=1, attention=False):
        self.C = C
        self.img_channels = img_channels
        self.attention = bool(attention)
        self.device = torch.device(device)
        self.specs = param_specs(C, img_channels, attention)
        self.offsets = {}
        off = 0
        # the fused emb_layers block first: weights [12*C][2C] contiguous, then biases [12*C]
        self.emb_w_off = off
        for rbname in RESBLOCKS:
            self.offsets[rbname + ".emb_layers.1.weight"] = off
            off += C * 2 * C
        self.emb_b_off = off
        for rbname in RESBLOCKS:
            self.offsets[rbname + ".emb_layers.1.bias"] = off
            off += C
        off = _align(off)
        for name, shape, _ in self.specs:
            if name in self.offsets:
                continue
            self.offsets[name] = off
            off = _align(off + math.prod(shape))
        self.numel = off
        # data-parallel bucket boundary: [bucket_split, numel) holds every out.* / up.* tensor (registration order puts
        # them last), whose gradients are complete before the backward reaches the turn block
        self.bucket_split = min(o for n, o in self.offsets.items() if n.startswith(("up.", "out.")) and ".emb_layers." not in n)
        assert all(o < self.bucket_split for n, o in self.offsets.items()
                   if not n.startswith(("up.", "out.")) or ".emb_layers." in n)
        self.flat = torch.zeros(off, dtype=torch.float32, device=self.device)
        self.grad = torch.zeros(off, dtype=torch.float32, device=self.device)
        # TF32-rounded (round-to-nearest) shadow of `flat`: what the tensor-core convolutions read as their B operand
        self.flat_tc = torch.zeros(off, dtype=torch.float32, device=self.device)
        self.flat_bf = None  # bf16 shadow (ensure_bf16): the operands of the bf16 ResBlock convolutions
        self.exp_avg = None
        self.exp_avg_sq = None
        self.shapes = {n: s for n, s, _ in self.specs}
        self.kinds = {n: k for n, _, k in self.specs}

    def _view(self, buf, name):
        shape = self.shapes[name]
        n = math.prod(shape)
        flat = buf[self.offsets[name]: self.offsets[name] + n]
        if self.kinds[name] == "conv_w":
            o, i, kh, kw = shape
            return flat.view(o, kh, kw, i).permute(0, 3, 1, 2)  # logical OIHW, physical [O][kh][kw][I]
        return flat.view(shape)

    def view(self, name):
        return self._view(self.flat, name)

    def grad_view(self, name):
        return self._view(self.grad, name)

    def ptr(self, name):
        return self.flat.data_ptr() + 4 * self.offsets[name]

    def tc_ptr(self, name):
        return self.flat_tc.data_ptr() + 4 * self.offsets[name]

    def ensure_bf16(self):
        if self.flat_bf is None:
            self.flat_bf = torch.zeros(self.numel, dtype=torch.bfloat16, device=self.device)

    def ensure_transposed(self, bf16):
        """Transposed, tap-flipped shadows of the conv weights ([Cin][T][128] at the weight's own offset): backward-data
        runs through the forward kernel with them.  Returns the device table {off, T, Cin, 0} of the 128-output convs."""
        if getattr(self, "flat_tcT", None) is None:
            self.flat_tcT = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
            rows = []
            for name, shape, kind in self.specs:
                if kind == "conv_w" and shape[0] == 128 and shape[1] % 32 == 0:
                    rows.append([self.offsets[name], shape[2] * shape[3], shape[1], 0])
            self.wT_table = torch.tensor(rows, dtype=torch.int32, device=self.device)
            self.wT_max = max(128 * r[1] * r[2] for r in rows)
        if bf16 and getattr(self, "flat_bfT", None) is None:
            self.flat_bfT = torch.zeros(self.numel, dtype=torch.bfloat16, device=self.device)
        return self.wT_table

    def tcT_ptr(self, name):
        return self.flat_tcT.data_ptr() + 4 * self.offsets[name]

    def bfT_ptr(self, name):
        return self.flat_bfT.data_ptr() + 2 * self.offsets[name]

    def bf_ptr(self, name):
        return self.flat_bf.data_ptr() + 2 * self.offsets[name]

    def gptr(self, name):
        return self.grad.data_ptr() + 4 * self.offsets[name]

    def init_like_reference(self):
        """Same distributions AND the same RNG consumption order as SimpleUnet.__init__ (nn.Linear / nn.Conv2d
        reset_parameters: kaiming_uniform(a=sqrt(5)) weight then uniform bias; GroupNorm ones/zeros; zero_module),
        so `torch.manual_seed(s)` gives the reference's initial weights bit for bit."""
        for name, shape, kind in self.specs:
            if kind in ("lin_w", "conv_w"):
                w = torch.empty(shape)
                nn.init.kaiming_uniform_(w, a=math.sqrt(5))
                fan_in = math.prod(shape[1:])
                bound = 1 / math.sqrt(fan_in) if fan_in > 0 else 0
                b = torch.empty(shape[0])
                nn.init.uniform_(b, -bound, bound)
                if is_zero_init(name):
                    w.zero_()
                    b.zero_()
                self.view(name).copy_(w)
                self.view(name[:-len("weight")] + "bias").copy_(b)
            elif kind == "gn_w":
                self.view(name).fill_(1.0)
            elif kind == "gn_b":
                self.view(name).zero_()

    def ensure_adam_state(self):
        if self.exp_avg is None:
            self.exp_avg = torch.zeros_like(self.flat)
            self.exp_avg_sq = torch.zeros_like(self.flat)


def build_module_tree(store, root):
    """Register every parameter on `root` (an nn.Module) under the reference's dotted names, creating plain
    nn.Module containers on the way, so state_dict() keys/shapes/order equal the reference's SimpleUnet."""
    for name, _, _ in store.specs:
        parts = name.split(".")
        mod = root
        for p in parts[:-1]:
            if p not in mod._modules:
                mod.add_module(p, nn.Module())
            mod = mod._modules[p]
        param = nn.Parameter(store.view(name), requires_grad=True)
        param.grad = store.grad_view(name)
        mod.register_parameter(parts[-1], param)
    return root
