"""ctypes binding of libpozmlp.so (the C ABI declared in include/pozmlp.h).

There is no CPU or PyTorch fallback: if the shared library is missing, or an entry point fails,
this module raises.  ``build()`` compiles the library in-tree for sm_100a.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libpozmlp.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "pozmlp.h")

ACT_CODES = {"none": 0, "relu": 1, "sigmoid": 2, "tanh": 3, "softmax": 4}
LOSS_MSE, LOSS_CE = 0, 1
DTYPE_F32, DTYPE_F64, DTYPE_BF16 = 0, 1, 2
OPT_SGD, OPT_ADAM = 0, 1

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]
SOURCES = ["poz_api.cu", "poz_noalloc.cu", "poz_ops.cu", "poz_tc.cu", "poz_tc3.cu"]


class PozError(RuntimeError):
    """A C-ABI call returned non-zero; the message is poz_last_error()."""


class LayerDesc(C.Structure):
    _fields_ = [("in_", C.c_int), ("out", C.c_int), ("act", C.c_int), ("has_bn", C.c_int),
                ("bn_eps", C.c_double), ("bn_momentum", C.c_double),
                ("drop_z", C.c_int), ("drop_bn", C.c_int), ("drop_a", C.c_int), ("row_mult", C.c_int)]


class StackDesc(C.Structure):
    _fields_ = [("n_linear", C.c_int), ("layers", C.POINTER(LayerDesc)),
                ("has_bias", C.c_int), ("loss_code", C.c_int), ("dtype_code", C.c_int),
                ("emb_vocab", C.c_int), ("emb_dim", C.c_int), ("emb_drop", C.c_int), ("tokens", C.c_int)]


class TrainConfig(C.Structure):
    _fields_ = [
        ("optimizer", C.c_int),
        ("learning_rate", C.c_double),
        ("decay_rate", C.c_double),
        ("dropout_rate", C.c_double),
        ("l2_lambda", C.c_double),
        ("beta1", C.c_double),
        ("beta2", C.c_double),
        ("epsilon", C.c_double),
        ("epochs", C.c_longlong),
        ("progress_stride", C.c_longlong),
        ("dropout_seed", C.c_ulonglong),
        ("global_batch", C.c_longlong),
    ]


def _sources():
    return [os.path.join(_CSRC, s) for s in SOURCES]


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    built = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC)] + [HEADER_PATH]
    return any(os.path.getmtime(d) > built for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile libpozmlp.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).  The translation
    units are compiled side by side (objects under build/obj, git-ignored) and linked into the one library;
    only units whose sources are newer than their object are recompiled."""
    if not force and not _stale():
        return LIB_PATH
    from concurrent.futures import ThreadPoolExecutor
    nvcc = os.environ.get("NVCC", "nvcc")
    obj_dir = os.path.join(os.path.dirname(_HERE), "build", "obj")
    os.makedirs(obj_dir, exist_ok=True)
    headers = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if not f.endswith(".cu")] + [HEADER_PATH]
    newest_header = max(os.path.getmtime(h) for h in headers)
    flags = [f for f in NVCC_FLAGS if f != "-shared"]

    def compile_one(src):
        obj = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".o")
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(src), newest_header):
            return obj, ""
        cmd = [nvcc] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
        return obj, proc.stderr

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as pool:
        results = list(pool.map(compile_one, _sources()))
    if verbose:
        print("".join(r[1] for r in results))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC"] + [r[0] for r in results] + ["-o", LIB_PATH]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    return LIB_PATH


_lib = None
_lock = threading.Lock()

_VP, _LL, _I, _DP = C.c_void_p, C.c_longlong, C.c_int, C.POINTER(C.c_double)

# name -> (restype, argtypes); mirrors include/pozmlp.h one to one
SIGNATURES = {
    "poz_last_error": (C.c_char_p, []),
    "poz_version": (_I, []),
    "poz_launch_count": (_LL, []),
    "poz_profile_begin": (_I, [_VP]),
    "poz_profile_end": (_I, [_VP, C.c_char_p, _LL]),
    "poz_ctx_create": (_I, [_I, _I, C.POINTER(_I), C.POINTER(_I), C.POINTER(C.c_ubyte),
                            C.POINTER(C.c_ubyte), _I, _I, _I, C.POINTER(_VP)]),
    "poz_ctx_create_ex": (_I, [_I, C.POINTER(StackDesc), C.POINTER(_VP)]),
    "poz_tensor_io": (_I, [_VP, _I, _I, _I, _I, _VP]),
    "poz_ctx_destroy": (_I, [_VP]),
    "poz_ctx_reserve": (_I, [_VP, _LL, _I]),
    "poz_params_upload": (_I, [_VP, _I, _VP, _VP]),
    "poz_params_download": (_I, [_VP, _I, _VP, _VP]),
    "poz_adam_upload": (_I, [_VP, _I, _VP, _VP, _VP, _VP]),
    "poz_adam_download": (_I, [_VP, _I, _VP, _VP, _VP, _VP]),
    "poz_adam_set_step": (_I, [_VP, _LL]),
    "poz_adam_get_step": (_I, [_VP, C.POINTER(_LL)]),
    "poz_forward": (_I, [_VP, _VP, _LL, _VP, _VP]),
    "poz_forward_loss": (_I, [_VP, _VP, _VP, _LL, _VP, _DP, _VP]),
    "poz_train_begin": (_I, [_VP, C.POINTER(TrainConfig)]),
    "poz_train_step": (_I, [_VP, _VP, _VP, _VP, _LL, C.POINTER(_VP), _I, _VP]),
    "poz_train_step_host": (_I, [_VP, _VP, _VP, _LL, _DP, _VP]),
    "poz_train_host_flush": (_I, [_VP, _DP]),
    "poz_train_costs": (_I, [_VP, _LL, _LL, _VP, _VP]),
    "poz_train_update_ratios": (_I, [_VP, _LL, _LL, _VP, _VP]),
    "poz_diag_activation": (_I, [_VP, _I, _I, _VP]),
    "poz_diag_activation_grad": (_I, [_VP, _I, _I, _VP, C.POINTER(_I)]),
    "poz_diag_weight_grad": (_I, [_VP, _I, _VP, _VP]),
    "poz_diag_moments": (_I, [_VP, _I, _I, _I, _I, C.c_double, _VP]),
    "poz_diag_histogram": (_I, [_VP, _I, _I, _I, _VP, _I, _VP]),
    "poz_grad_buffer": (_I, [_VP, C.POINTER(_VP), C.POINTER(_LL), C.POINTER(_I), _VP]),
    "poz_grad_layer": (_I, [_VP, _I, C.POINTER(_VP), C.POINTER(_LL), _VP]),
    "poz_train_backward": (_I, [_VP, _VP, _VP, _VP, _LL, C.POINTER(_VP), _I, _VP]),
    "poz_train_apply": (_I, [_VP, _VP]),
    "poz_op_linear": (_I, [_I, _I, _VP, _VP, _VP, _LL, _I, _I, _VP, _VP]),
    "poz_op_activation": (_I, [_I, _I, _I, _VP, _LL, _I, _VP, _VP]),
    "poz_op_embedding": (_I, [_I, _I, _VP, _I, _I, _VP, _LL, _VP, _VP]),
    "poz_op_batchnorm": (_I, [_I, _I, _I, _I, _VP, _LL, _I, _VP, _VP, _VP, _VP, C.c_double, C.c_double, _VP, _VP]),
    "poz_tc_gemm": (_I, [_I, _I, _VP, _VP, _LL, _I, _LL, _VP, _VP, _VP, _I, _I, _VP]),
    "poz_tc3_gemm": (_I, [_I, _I, _VP, _VP, _LL, _I, _LL, _VP, _VP, _VP, _I, _I, _VP]),
    "poz_forward_noalloc": (_I, [_I, _I, C.POINTER(_I), C.POINTER(_I), C.POINTER(_VP), C.POINTER(_VP),
                                _VP, _LL, _VP, _VP]),
}
NOALLOC_MAX_WIDTH = 4096
NOALLOC_SCRATCH_BYTES = 8 << 20


def lib():
    """The loaded library.  Raises if it has not been built -- there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise PozError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                               "g.build()'` (there is no CPU fallback for the hot path)")
            handle = C.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(handle, name)      # AttributeError if the symbol is not exported
                fn.restype = res
                fn.argtypes = args
            _lib = handle
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().poz_last_error()
        raise PozError(msg.decode() if msg else f"pozmlp call failed with code {rc}")


def launch_count() -> int:
    return int(lib().poz_launch_count())
