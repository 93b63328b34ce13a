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
    ]


class WgradArgs(C.Structure):
    _fields_ = [
        ("x", _f32p), ("x2", _f32p), ("dy", _f32p), ("dw", _f32p),
        ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("C1", C.c_int), ("C2", C.c_int),
        ("Ho", C.c_int), ("Wo", C.c_int),
        ("R", C.c_int), ("S", C.c_int), ("stride", C.c_int), ("pad", C.c_int),
        ("k_slices", C.c_int),
    ]


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
    if lib.b200_abi_version() != 1:
        raise B200Error("libb200ddpm.so ABI version mismatch")
    for name in EXPORTS:
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        if name.startswith("b200_") and name not in ("b200_abi_version", "b200_init", "b200_last_error",
                                                       "b200_launch_count"):
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    _lib = lib
    return lib


# every symbol include/b200ddpm.h declares (tests check the .so exports each of them)
EXPORTS = [
    "b200_abi_version", "b200_init", "b200_last_error", "b200_launch_count",
    "b200_conv_igemm", "b200_conv_wgrad",
]


def call(name, args, stream, ws=None, ws_bytes=0):
    """Invoke an `int fn(const Args*, void* ws, size_t ws_bytes, void* stream)` entry point."""
    lib = load()
    rc = getattr(lib, name)(C.byref(args), ws, ws_bytes, stream)
    if rc != 0:
        raise B200Error(f"{name} failed (rc={rc}): {lib.b200_last_error().decode()}")


def launch_count():
    return int(load().b200_launch_count())
