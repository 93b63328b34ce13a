"""ctypes binding of libb200ddpm.so (the C ABI declared in include/b200ddpm.h).

There is deliberately no fallback: if the shared library is missing or a call is rejected, this module raises.
The library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a) and lives next to this file.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ddpm.so")

_f32p = C.c_void_p  # raw device pointers travel as void*


class ConvArgs(C.Structure):
    _fields_ = [
        ("x", _f32p), ("x2", _f32p), ("w", _f32p), ("out", _f32p),
        ("bias", _f32p), ("emb", _f32p), ("residual", _f32p),
        ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("C1", C.c_int), ("C2", C.c_int),
        ("Ho", C.c_int), ("Wo", C.c_int),
        ("R", C.c_int), ("S", C.c_int), ("stride", C.c_int), ("pad", C.c_int),
        ("mode", C.c_int), ("w_cout", C.c_int), ("w_cin", C.c_int), ("n_off", C.c_int),
        ("out_ld", C.c_int), ("emb_ld", C.c_int), ("res_ld", C.c_int),
        ("x_ld", C.c_int), ("x2_ld", C.c_int), ("round_out", C.c_int), ("out_dtype", C.c_int),
        ("skip_x", _f32p), ("skip_x2", _f32p), ("skip_w", _f32p), ("bias2", _f32p),
        ("skip_c1", C.c_int), ("skip_c2", C.c_int), ("skip_x_ld", C.c_int), ("skip_x2_ld", C.c_int),
        ("op_dtype", C.c_int), ("out_bf16", _f32p),
    ]


class WgradArgs(C.Structure):
    _fields_ = [
        ("x", _f32p), ("x2", _f32p), ("dy", _f32p), ("dw", _f32p),
        ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("C1", C.c_int), ("C2", C.c_int),
        ("Ho", C.c_int), ("Wo", C.c_int),
        ("R", C.c_int), ("S", C.c_int), ("stride", C.c_int), ("pad", C.c_int),
        ("k_slices", C.c_int), ("x_ld", C.c_int), ("x2_ld", C.c_int), ("dy_ld", C.c_int), ("op_dtype", C.c_int),
    ]


class GnArgs(C.Structure):
    _fields_ = [
        ("x", _f32p), ("x2", _f32p), ("stats", _f32p), ("gamma", _f32p), ("beta", _f32p), ("y", _f32p),
        ("N", C.c_int), ("HW", C.c_int), ("C1", C.c_int), ("C2", C.c_int),
        ("x_ld", C.c_int), ("x2_ld", C.c_int),
        ("eps", C.c_float), ("silu", C.c_int), ("round_tf32", C.c_int), ("y_dtype", C.c_int), ("x_dtype", C.c_int),
        ("dropout_p", C.c_float), ("dropout_seed", C.c_uint), ("dropout_step", _f32p),
    ]


class GnBwdArgs(C.Structure):
    _fields_ = [
        ("fwd", GnArgs), ("dy", _f32p), ("dx", _f32p), ("add1", _f32p), ("add2", _f32p),
        ("add1_ld", C.c_int), ("add2_ld", C.c_int), ("add_c", C.c_int),
        ("dgamma", _f32p), ("dbeta", _f32p), ("round_dx", C.c_int), ("dx_bf16", _f32p), ("dx_bf16_c", C.c_int),
        ("colsum_per_sample", _f32p), ("colsum_ps_ld", C.c_int), ("colsum_total1", _f32p), ("colsum_total2", _f32p),
        ("colsum_c", C.c_int), ("dy_dtype", C.c_int),
    ]


class QSampleArgs(C.Structure):
    _fields_ = [("x", _f32p), ("eps", _f32p), ("logsnr", _f32p), ("z", _f32p), ("B", C.c_int), ("D", C.c_int)]


class LossArgs(C.Structure):
    _fields_ = [
        ("out", _f32p), ("z", _f32p), ("x", _f32p), ("eps", _f32p), ("logsnr", _f32p),
        ("loss_b", _f32p), ("mse2", _f32p), ("dout", _f32p), ("loss_mean", _f32p),
        ("B", C.c_int), ("D", C.c_int), ("mean_type", C.c_int), ("grad_scale", C.c_float), ("loss_type", C.c_int),
    ]


class SamplerStepArgs(C.Structure):
    _fields_ = [
        ("z", _f32p), ("out_c", _f32p), ("out_u", _f32p), ("cond_w", _f32p), ("noise", _f32p), ("sched", _f32p),
        ("step_dev", _f32p), ("z_next", _f32p), ("x_hist", _f32p), ("e_hist", _f32p),
        ("n", C.c_int), ("D", C.c_int), ("step", C.c_int), ("mean_type", C.c_int), ("sampler", C.c_int),
        ("z_dup", _f32p), ("z_hist", _f32p), ("hist_T", C.c_int), ("z_u8", _f32p), ("x_u8", _f32p), ("e_u8", _f32p),
        ("logsnr_t_b", _f32p), ("logsnr_s_b", _f32p),
    ]


class DistillTargetArgs(C.Structure):
    _fields_ = [
        ("z_teacher", _f32p), ("z_t", _f32p), ("x_pred_teacher", _f32p), ("logsnr_t", _f32p), ("logsnr_s", _f32p),
        ("index", _f32p), ("x_target", _f32p), ("eps_target", _f32p), ("n", C.c_int), ("D", C.c_int),
    ]


class StepAdvanceArgs(C.Structure):
    _fields_ = [("step_dev", _f32p)]


class ConvInArgs(C.Structure):
    _fields_ = [
        ("x", _f32p), ("w", _f32p), ("bias", _f32p), ("out", _f32p), ("dy", _f32p), ("dw", _f32p), ("db", _f32p),
        ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("round_out", C.c_int),
        ("x_ld", C.c_int), ("w_co_ld", C.c_int), ("w_tap_ld", C.c_int), ("accumulate", C.c_int),
    ]


class ConvOutArgs(C.Structure):
    _fields_ = [
        ("h", _f32p), ("w", _f32p), ("bias", _f32p), ("out", _f32p), ("dout", _f32p), ("dh", _f32p),
        ("dw", _f32p), ("db", _f32p), ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("round_dh", C.c_int),
        ("out_ld", C.c_int), ("accumulate_dh", C.c_int),
    ]


class ResampleArgs(C.Structure):
    _fields_ = [("in_", _f32p), ("out", _f32p), ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("C", C.c_int),
                ("in_ld", C.c_int), ("mode", C.c_int), ("round_out", C.c_int), ("out_dtype", C.c_int)]


class ColsumArgs(C.Structure):
    _fields_ = [("dy", _f32p), ("per_sample", _f32p), ("total1", _f32p), ("total2", _f32p),
                ("N", C.c_int), ("HW", C.c_int), ("ld", C.c_int), ("ps_ld", C.c_int)]


class EmbedArgs(C.Structure):
    _fields_ = [
        ("t", _f32p), ("sched", _f32p), ("step_dev", _f32p), ("freqs", _f32p), ("guide", _f32p),
        ("sincos", _f32p), ("onehot", _f32p), ("t_out", _f32p), ("t_emb", _f32p), ("g_emb", _f32p),
        ("emb", _f32p), ("semb", _f32p), ("d_semb", _f32p), ("d_emb", _f32p), ("d_g_emb", _f32p),
        ("B", C.c_int), ("E", C.c_int), ("step", C.c_int), ("force_uncond", C.c_int), ("w_emb", _f32p),
    ]


class LinearArgs(C.Structure):
    _fields_ = [
        ("X", _f32p), ("W", _f32p), ("bias", _f32p), ("Y", _f32p), ("dY", _f32p), ("dX", _f32p), ("dW", _f32p),
        ("db", _f32p), ("M", C.c_int), ("N", C.c_int), ("K", C.c_int), ("act", C.c_int), ("mode", C.c_int),
    ]


class AdamArgs(C.Structure):
    _fields_ = [
        ("p", _f32p), ("g", _f32p), ("m", _f32p), ("v", _f32p), ("n", C.c_size_t),
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("grad_scale", C.c_float),
        ("step_dev", _f32p), ("step", C.c_int), ("hparams_dev", _f32p),
    ]


class RoundArgs(C.Structure):
    _fields_ = [("src", _f32p), ("dst", _f32p), ("n", C.c_size_t)]


class CastArgs(C.Structure):
    _fields_ = [("src", _f32p), ("dst", _f32p), ("rows", C.c_size_t), ("cols", C.c_int), ("src_ld", C.c_int)]


class WTransposeArgs(C.Structure):
    _fields_ = [("src", _f32p), ("dst", _f32p), ("table", _f32p), ("n_weights", C.c_int), ("max_elems", C.c_int),
                ("out_dtype", C.c_int)]


class ZeroArgs(C.Structure):
    _fields_ = [("ptr", _f32p), ("bytes", C.c_size_t)]


class BgemmArgs(C.Structure):
    _fields_ = [
        ("A", _f32p), ("B", _f32p), ("C", _f32p), ("M", C.c_int), ("N", C.c_int), ("K", C.c_int), ("batch", C.c_int),
        ("lda", C.c_int), ("ldb", C.c_int), ("ldc", C.c_int),
        ("strideA", C.c_longlong), ("strideB", C.c_longlong), ("strideC", C.c_longlong),
        ("a_mn", C.c_int), ("b_mn", C.c_int), ("alpha", C.c_float), ("round_out", C.c_int),
    ]


class SoftmaxArgs(C.Structure):
    _fields_ = [("s", _f32p), ("p", _f32p), ("dp", _f32p), ("ds", _f32p), ("rows", C.c_longlong), ("T", C.c_int),
                ("ld", C.c_int)]


class B200Error(RuntimeError):
    pass


_lib = None


def load():
    """Load libb200ddpm.so once; raise (never fall back) if it is absent or has the wrong ABI."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU or PyTorch fallback for the DDPM hot path."
        )
    lib = C.CDLL(LIB_PATH)
    lib.b200_abi_version.restype = C.c_int
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_launch_count.restype = C.c_ulonglong
    lib.b200_init.argtypes = [C.c_int]
    if lib.b200_abi_version() != 6:
        raise B200Error("libb200ddpm.so ABI version mismatch")
    for name in EXPORTS:
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        if name == "b200_conv_kernel_of":
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p]
        elif name == "b200_gn_bwd_workspace":
            fn.restype = C.c_size_t
            fn.argtypes = [C.c_void_p]
        elif name == "b200_gn_bwd_static":  # (args, stream): the scratch is a static device array inside the library
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p, C.c_void_p]
        elif name.startswith("b200_") and name not in ("b200_abi_version", "b200_init", "b200_last_error",
                                                         "b200_launch_count"):
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    _lib = lib
    return lib


# every symbol include/b200ddpm.h declares (tests check the .so exports each of them)
EXPORTS = [
    "b200_abi_version", "b200_init", "b200_last_error", "b200_launch_count",
    "b200_conv_igemm", "b200_conv_kernel_of", "b200_conv_wgrad",
    "b200_gn_fwd", "b200_gn_bwd", "b200_gn_bwd_workspace", "b200_gn_bwd_static",
    "b200_q_sample", "b200_loss", "b200_sampler_step", "b200_distill_target", "b200_step_advance", "b200_int_increment",
    "b200_bgemm", "b200_softmax_fwd", "b200_softmax_bwd",
    "b200_conv_in_fwd", "b200_conv_in_bwd", "b200_conv_out_fwd", "b200_conv_out_bwd",
    "b200_upsample2x", "b200_colsum",
    "b200_embed_prep", "b200_embed_combine", "b200_embed_combine_bwd",
    "b200_linear", "b200_adam", "b200_zero", "b200_round_tf32", "b200_cast_bf16", "b200_conv_weights_transpose",
]


def call(name, args, stream, ws=None, ws_bytes=0):
    """Invoke an `int fn(const Args*, void* ws, size_t ws_bytes, void* stream)` entry point."""
    lib = load()
    rc = getattr(lib, name)(C.byref(args), ws, ws_bytes, stream)
    if rc != 0:
        raise B200Error(f"{name} failed (rc={rc}): {lib.b200_last_error().decode()}")


def launch_count():
    return int(load().b200_launch_count())
