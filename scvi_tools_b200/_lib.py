"""ctypes binding of the C-ABI library ``libscvi_b200.so`` (``include/scvi_b200.h``).

There is no fallback: if the library is missing or a symbol is absent, importing the ops
raises.  Every call goes through :func:`call`, which maps non-zero return codes to
``RuntimeError`` carrying ``scvi_last_error_string()``.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libscvi_b200.so")

P = C.c_void_p
I64 = C.c_int64
I32 = C.c_int
F32 = C.c_float

# name -> (restype, argtypes); mirrors include/scvi_b200.h declaration by declaration
SIGNATURES = {
    "scvi_b200_abi_version": (C.c_int, []),
    "scvi_last_error_string": (C.c_char_p, []),
    "scvi_launch_count": (C.c_int64, []),
    "scvi_profile_begin": (C.c_int, [P, I64]),
    "scvi_profile_end": (C.c_int64, [P, I64]),
    "scvi_gather_rows": (C.c_int, [P, I64, I64, I64, P, I64, P, I64, P]),
    "scvi_gather_i64": (C.c_int, [P, I64, P, I64, P, P]),
    "scvi_gather_rows_u16": (C.c_int, [P, I64, I64, I64, P, I64, P, I64, P]),
    "scvi_gather_rows_u8": (C.c_int, [P, I64, I64, I64, P, I64, P, I64, P]),
    "scvi_host_gather_rows": (C.c_int, [P, I64, I64, I64, P, I64, P, I64, C.c_int]),
    "scvi_gather_rows_u8_raw": (C.c_int, [P, I64, I64, I64, P, I64, P, I64, P]),
    "scvi_gather_rows_csr": (C.c_int, [P, P, C.c_int, P, C.c_int, I64, I64, P, I64, P, I64, P, P]),
    "scvi_nb_logprob_fwd": (C.c_int, [P, I64, P, I64, P, I64, I64, I64, F32, P, I64, P, P]),
    "scvi_zinb_logprob_fwd": (C.c_int, [P, I64, P, I64, P, I64, P, I64, I64, I64, F32, P, I64, P, P]),
    "scvi_nb_logprob_bwd": (C.c_int, [P, I64, P, I64, P, I64, P, I64, I64, I64, F32, P, P, P]),
    "scvi_zinb_logprob_bwd": (C.c_int, [P, I64, P, I64, P, I64, P, I64, P, I64, I64, I64, F32, P, P, P, P]),
    "scvi_mixture_nb_logprob_fwd": (C.c_int, [P, I64, P, I64, P, I64, P, I64, P, I64, I64, I64, F32, P, I64, P, P]),
    "scvi_mixture_nb_logprob_bwd": (
        C.c_int, [P, I64, P, I64, P, I64, P, I64, P, I64, P, I64, I64, I64, F32, P, P, P, P, P]),
    "scvi_colsum_partials": (C.c_int, [P, I64, I64, P, I32, P]),
    "scvi_linear_fwd_ws_bytes": (I64, [I64, I64, I64, I32]),
    "scvi_linear_fwd": (C.c_int, [P, I64, P, I64, I64, I64, P, I64, P, P, I32, I32, P, I64, P, P, P, P]),
    "scvi_linear_bwd_ws_bytes": (I64, [I64, I64, I64, I64]),
    "scvi_linear_bwd": (C.c_int, [P, I64, P, I64, I64, I64, P, I64, I64, P, I32, I32, P, I64, P, P, P, I64, P, I64, P, P]),
    "scvi_bn_finalize": (C.c_int, [P, I64, I64, I32, F32, F32, P, P, P, P, P]),
    "scvi_bn_partials": (C.c_int, [P, I64, I64, I64, P, P]),
    "scvi_bn_act_drop_fwd": (C.c_int, [P, I64, I64, I64, P, P, P, P, I32, P, F32, P, I64, P]),
    "scvi_bn_bwd_ws_bytes": (I64, [I64, I64]),
    "scvi_bn_act_drop_bwd": (C.c_int, [P, I64, P, I64, I64, I64, P, P, P, P, I32, P, F32, I32, P, I64, P, P, P, P]),
    "scvi_bn_static_scratch_bytes": (I64, []),
    "scvi_bn_act_drop_bwd_static": (C.c_int, [P, I64, P, I64, I64, I64, P, P, P, P, I32, P, F32, I32, P, I64, P, P, P]),
    "scvi_ln_act_drop_fwd": (C.c_int, [P, I64, I64, I64, F32, I32, P, F32, P, I64, P, P]),
    "scvi_ln_act_drop_bwd": (C.c_int, [P, I64, P, I64, I64, I64, P, I32, P, F32, P, I64, P]),
    "scvi_embedding_bwd": (C.c_int, [P, I64, P, I64, I64, I64, P, P]),
    "scvi_latent_fwd": (C.c_int, [P, I64, I64, I64, I64, P, P, P, P, P, F32, P, P, P, P, P]),
    "scvi_latent_bwd_ws_bytes": (I64, [I64, I64, I64]),
    "scvi_latent_bwd": (C.c_int, [P, I64, I64, I64, I64, P, P, P, F32, P, P, P, P, P, P, P, I64, P, P, P, P, P, P]),
    "scvi_mid_ws_bytes": (I64, [I64, I64, I64]),
    "scvi_mid_fwd": (C.c_int, [P, P, I64, I64, I64, P, P, P, P, P, F32, P, P, P, P, P, F32, P, I64, P, P, P, P, P, P,
                               F32, F32, P, P, P, P, P, P, P, P, P, P, P, P, P]),
    "scvi_mid_bwd": (C.c_int, [P, P, P, P, P, P, P, P, I64, P, P, P, P, P, F32, P, P, P, P, P, P, P, P, P, F32, I64, I64,
                               I64, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P]),
    "scvi_mid_n_tiles": (I64, [I64]),
    "scvi_decoder_nll_fwd_ws_bytes": (I64, [I64, I64, I64, I32, I32]),
    "scvi_decoder_nll_fwd": (C.c_int, [P, I64, I64, I64, I64, P, P, P, P, P, P, I64, P, P, I32, I32, P, P, P, P, I64, P, P]),
    "scvi_decoder_nll_prepack": (C.c_int, [I64, I64, I64, P, P, P, P, P, I32, I32, I32, P, P]),
    "scvi_decoder_nll_bwd_ws_bytes": (I64, [I64, I64, I64, I32]),
    "scvi_decoder_nll_bwd": (
        C.c_int,
        [P, I64, I64, I64, I64, P, P, P, P, P, P, I64, P, P, P, P, I32, I32, P, I64, P, P, P, P, P, P, P, P],
    ),
    "scvi_decoder_params_fwd": (C.c_int, [P, I64, I64, I64, I64, P, P, P, P, P, P, P, P, I64, P]),
    "scvi_adam_step": (C.c_int, [P, P, P, P, I64, I64, F32, F32, F32, F32, F32, F32, P]),
    "scvi_normal_logprob_sum_fwd": (C.c_int, [P, P, P, I64, I64, I32, P, P]),
    "scvi_normal_logprob_sum_bwd": (C.c_int, [P, P, P, P, I64, I64, I32, P, P, P, P]),
    "scvi_categorical_kl": (C.c_int, [P, P, I64, I64, I64, P, P, P, P]),
    "scvi_gather_segments": (C.c_int, [P, P, P, I32, P, P]),
    "scvi_adam_step_graph": (C.c_int, [P, P, P, P, I64, P, F32, F32, F32, F32, F32, F32, P]),
    "scvi_host_device_pointer": (C.c_int, [P, C.POINTER(C.c_void_p)]),
    "scvi_comm_ctrl_bytes": (I64, []),
    "scvi_comm_alloc": (C.c_int, [I64, C.POINTER(C.c_void_p), P]),
    "scvi_comm_open": (C.c_int, [P, C.POINTER(C.c_void_p)]),
    "scvi_comm_close": (C.c_int, [P]),
    "scvi_comm_free": (C.c_int, [P]),
    "scvi_allreduce_adam": (C.c_int, [P, P, I32, I32, I32, P, P, P, P, I64, I64, P, I32, F32, F32, F32, F32, F32, I32, P]),
    "scvi_allreduce_adam_sharded": (C.c_int, [P, P, P, I32, I32, I32, P, P, I64, P, F32, F32, F32, F32, F32, I32, P]),
    "scvi_elbo_reduce": (C.c_int, [P, P, I64, F32, P, P, P]),
    "scvi_elbo_seed": (C.c_int, [I64, F32, P, F32, P, P, P]),
    "scvi_step_prologue": (C.c_int, [C.c_uint64, P, I64, I64, I64, I64, F32, P, P, P, P, P, F32, P, F32, P, P, P, P, P]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C scvi_tools_b200/csrc`.  scvi_tools_b200 has no CPU / PyTorch fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def call(name: str, *args):
    """Call a C-ABI entry point; raise RuntimeError on a non-zero status."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.scvi_last_error_string().decode()
        raise RuntimeError(f"{name} failed (code {rc}): {msg}")


def query(name: str, *args) -> int:
    return int(getattr(load(), name)(*args))


# constants of include/scvi_b200.h
BN_TILE = 64
LIK_NB, LIK_ZINB = 0, 1
LIN_LOG1P, LIN_LIBRARY, LIN_TC = 1, 2, 4
PREC_AOPS = 0x800  # decoder fwd keeps its operands, bwd is a pure GEMM (integer count kinds)
LIN_X_U8 = 0x400  # scvi_linear_fwd / _bwd: uint8 counts (with LIN_TC)
PREC_FP32, PREC_TC = 0, 1
PREC_PREPACKED = 0x100
X_INTEGER, X_U8 = 0x200, 0x400  # count-operand kind OR-ed into `precision` (SCVI_X_INTEGER / SCVI_X_U8)
