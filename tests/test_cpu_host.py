"""CPU-side checks: the C-ABI library builds, loads and exports every symbol include/b200ddpm.h declares; the host
logic (flat parameter store, state_dict contract, schedule table) matches the reference's conventions; the product
path refuses to run without CUDA (no fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

from oracle import ddpm_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(built_lib):
    hdr = open(os.path.join(ROOT, "include", "b200ddpm.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(built_lib.EXPORTS), declared ^ set(built_lib.EXPORTS)
    so = C.CDLL(built_lib.LIB_PATH)
    for name in declared:
        assert hasattr(so, name), name
    assert built_lib.load().b200_abi_version() == 6


def test_ctypes_structs_match_header_field_order(built_lib):
    """every struct in the header has a ctypes mirror with the same field names in the same order."""
    hdr = open(os.path.join(ROOT, "include", "b200ddpm.h")).read()
    mirror = {"b200_conv_args": built_lib.ConvArgs, "b200_wgrad_args": built_lib.WgradArgs, "b200_gn_args": built_lib.GnArgs,
              "b200_gn_bwd_args": built_lib.GnBwdArgs, "b200_q_sample_args": built_lib.QSampleArgs,
              "b200_loss_args": built_lib.LossArgs, "b200_sampler_step_args": built_lib.SamplerStepArgs,
              "b200_step_advance_args": built_lib.StepAdvanceArgs, "b200_conv_in_args": built_lib.ConvInArgs,
              "b200_conv_out_args": built_lib.ConvOutArgs, "b200_resample_args": built_lib.ResampleArgs,
              "b200_colsum_args": built_lib.ColsumArgs, "b200_embed_args": built_lib.EmbedArgs,
              "b200_linear_args": built_lib.LinearArgs, "b200_zero_args": built_lib.ZeroArgs,
              "b200_round_args": built_lib.RoundArgs, "b200_cast_args": built_lib.CastArgs, "b200_wtranspose_args": built_lib.WTransposeArgs,
              "b200_adam_args": built_lib.AdamArgs, "b200_distill_target_args": built_lib.DistillTargetArgs, "b200_bgemm_args": built_lib.BgemmArgs,
              "b200_softmax_args": built_lib.SoftmaxArgs}
    for m in re.finditer(r"typedef struct \{(.*?)\}\s*(b200_\w+);", hdr, re.S):
        body, name = m.group(1), m.group(2)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        fields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            names = decl.split(",")
            first = names[0].split()[-1]
            fields.append(first.lstrip("*"))
            fields += [n.strip().lstrip("*") for n in names[1:]]
        got = [f[0].rstrip("_") for f in mirror[name]._fields_]
        assert got == fields, (name, got, fields)


def test_model_refuses_cpu(built_lib):
    from generative_models_b200.model import DiffusionModelB200, default_G
    with pytest.raises(built_lib.B200Error):
        DiffusionModelB200(default_G(device="cpu"))


def test_param_store_layout_and_state_dict_contract():
    from torch import nn
    from generative_models_b200.unet_spec import ParamStore, build_module_tree, param_specs
    st = ParamStore(128, "cpu")
    assert [(n, s) for n, s, _ in param_specs(128)] == O.param_shapes(128)
    net = build_module_tree(st, nn.Module())
    sd = net.state_dict()
    assert [(k, tuple(v.shape)) for k, v in sd.items()] == O.param_shapes(128)
    assert sum(v.numel() for v in sd.values()) == 6_033_665
    # conv weights are logical OIHW over a physical [O][kh][kw][I] block of the flat buffer
    w = st.view("up.seq.4.in_layers.2.weight")
    assert tuple(w.shape) == (128, 256, 3, 3) and w.stride() == (9 * 256, 1, 3 * 256, 256)
    ref = O.random_state_dict(3)
    net.load_state_dict(ref)
    off = st.offsets["up.seq.4.in_layers.2.weight"]
    phys = st.flat[off: off + 128 * 9 * 256].view(128, 3, 3, 256)
    assert torch.equal(phys, ref["up.seq.4.in_layers.2.weight"].permute(0, 2, 3, 1))
    # the 12 emb_layers weights are one contiguous [1536, 256] block in execution order
    from generative_models_b200.unet_spec import RESBLOCKS
    blk = st.flat[st.emb_w_off: st.emb_w_off + 1536 * 256].view(12, 128, 256)
    for i, rb in enumerate(RESBLOCKS):
        assert torch.equal(blk[i], ref[rb + ".emb_layers.1.weight"])
    # gradients are views of the flat gradient buffer (the allreduce bucket)
    p = dict(net.named_parameters())["out.2.bias"]
    st.grad[st.offsets["out.2.bias"]] = 7.0
    assert p.grad.item() == 7.0
    # no two tensors overlap
    spans = sorted((st.offsets[n], st.offsets[n] + torch.Size(s).numel()) for n, s, _ in st.specs)
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:])) and spans[-1][1] <= st.numel


def test_init_like_reference_consumes_rng_like_simple_unet():
    """Same seed => same initial weights as the reference's SimpleUnet(128, 0.0) constructor; pinned by a fixture
    checksum generated from the reference (oracle/make_golden.py, key init_*) when present, else by distribution."""
    from generative_models_b200.unet_spec import ParamStore
    torch.manual_seed(0)
    st = ParamStore(128, "cpu")
    st.init_like_reference()
    assert st.view("turn.out_layers.3.weight").abs().max().item() == 0
    assert st.view("out.0.weight").eq(1).all() and st.view("out.0.bias").eq(0).all()
    w = st.view("time_embed.0.weight")
    assert w.abs().max().item() <= 1 / 8 + 1e-7  # U(-1/sqrt(64), 1/sqrt(64))


def test_init_bit_exact_vs_reference_fixture(golden):
    if "init_seed0_checks" not in golden:
        pytest.skip("fixture predates the init check")
    from generative_models_b200.unet_spec import ParamStore
    torch.manual_seed(0)
    st = ParamStore(128, "cpu")
    st.init_like_reference()
    for k, v in golden["init_seed0_checks"].items():
        off = st.offsets[k]  # physical order: conv weights are [O][kh][kw][I], as the fixture stores them
        assert torch.equal(st.flat[off: off + v.numel()], v), k


def test_sampler_table_matches_scalar_form_and_fixture(golden):
    from generative_models_b200.schedule import logsnr_schedule_cosine, sampler_table, sampler_table_scalar
    for T in (1, 8, 250, 1000):
        assert torch.equal(sampler_table(T, "cpu"), sampler_table_scalar(T, "cpu"))
    assert torch.equal(sampler_table(8, "cpu"), golden["sched_T8"])
    assert torch.equal(logsnr_schedule_cosine(golden["sched_u"]), golden["sched_logsnr"])


def test_dg_keys_and_types_match_reference_contract():
    from pathlib import Path
    from generative_models_b200.model import DiffusionModelB200 as M
    want = dict(binarize=int, timesteps=int, hidden_size=int, dropout=float, sampler=str, mean_type=str, eval_heavy=int,
                class_cond=int, sample_cond_w=float, cf_drop_prob=float, teacher_path=type(Path(".")), teacher_mode=str,
                lr_scheduler=str)
    for k, t in want.items():
        assert k in M.DG and type(M.DG[k]) is t, k
    assert (M.DG.timesteps, M.DG.sampler, M.DG.mean_type, M.DG.hidden_size) == (250, "ddim", "v", 128)


def test_launcher_registers_model_in_reference_registry():
    """tools/run_gms_main.py: the unmodified registry lookup of gms/main.py:64,66 resolves to the B200 model, and the
    synthetic loader stands in for common.load_mnist (needs the reference tree: build container only)."""
    import importlib.util
    import os

    import pytest

    if not os.path.isdir("/root/reference/gms"):
        pytest.skip("reference tree absent (GPU box)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("run_gms_main", os.path.join(root, "tools", "run_gms_main.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    common = mod.install("/root/reference", synthetic=True)
    from generative_models_b200.model import DiffusionModelB200

    models = common.discover_models()
    assert models["diffusion_model"] is DiffusionModelB200
    assert "vae" in models  # the stock registry is still there
    train, test = common.load_mnist(8, binarize=False)
    x, y = next(iter(train))
    assert x.shape == (8, 1, 28, 28) and x.min() >= -1 and x.max() <= 1 and y.dtype == torch.int64


@pytest.mark.parametrize("mode", ["reference", "port"])
def test_cpu_reference_arm_emits_the_bench_schema(mode):
    """oracle/cpu_reference_bench.py — what `bench.py --impl reference` and the `cpu_baseline` leg run: the unmodified
    reference through gms.main's own construction path (when tools/vendor_reference.sh has installed it into baseline/_ref),
    else the oracle port.  Tiny sizes: this checks the plumbing, not a rate."""
    import json
    import subprocess
    import sys
    from oracle import reference_runtime as R
    if mode == "reference" and not R.available():
        pytest.skip("reference not vendored on this machine")
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "cpu_reference_bench.py"), "--steps", "1", "--warmup", "0", "--bs", "4",
           "--sample-n", "2", "--sample-steps", "1", "--threads", "4"] + (["--port"] if mode == "port" else [])
    r = subprocess.run(cmd, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""), stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    res = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert res["train"]["img_s"] > 0 and res["train"]["batch"] == 4 and res["threads"] == 4
    # (one cold step each: no ordering between the two rates is asserted, only that both legs ran)
    assert res["sampling"]["unguided"]["img_s_1000_steps"] > 0 and res["sampling"]["guided"]["img_s_1000_steps"] > 0
    assert res.get("kind", "reference") == mode and res["cuda_visible"] is False
