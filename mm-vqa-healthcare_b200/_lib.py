"""ctypes binding of libm3ae_b200.so (the C-ABI declared in include/m3ae_b200.h).

The product path has no CPU or PyTorch fallback: if the shared library is missing, or a call returns a
non-zero status, this module raises.  Nothing here imports or calls `oracle/`.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# M3AE_B200_LIB: developer hook for same-box A/B timing of two builds of the library (tools/ab/); unset in every test and bench record
LIB_PATH = os.environ.get("M3AE_B200_LIB") or os.path.join(_HERE, "libm3ae_b200.so")

F32, BF16 = 0, 1
ACT_NONE, ACT_GELU, ACT_RELU, ACT_QUICKGELU, ACT_TANH = 0, 1, 2, 3, 4
MAJOR_K, MAJOR_MN = 0, 1
GEMM_AUTO, GEMM_TCGEN05, GEMM_SIMT = 0, 1, 2
ATTN_MMA_SYNC = 3
ATTN_NARROW = 4


class GemmEpilogue(C.Structure):
    _fields_ = [
        ("out", C.c_void_p), ("out_dtype", C.c_int32), ("ldo", C.c_int64),
        ("bias", C.c_void_p), ("act", C.c_int32),
        ("aux_out", C.c_void_p), ("aux_in", C.c_void_p), ("aux_in_dtype", C.c_int32), ("ld_aux_in", C.c_int64),
        ("dact", C.c_int32),
        ("residual", C.c_void_p), ("res_dtype", C.c_int32), ("ldr", C.c_int64),
        ("drop_p", C.c_float), ("drop_seed", C.c_uint64), ("drop_stream", C.c_uint64),
        ("accumulate", C.c_int32), ("alpha", C.c_float), ("aux_is_grad", C.c_int32),
        ("drop_bits", C.c_void_p), ("drop_bits_ready", C.c_int32),
        ("amax_val", C.c_void_p), ("amax_idx", C.c_void_p),
    ]


class GemmProblem(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("b", C.c_void_p), ("lda", C.c_int64), ("ldb", C.c_int64),
        ("a_major", C.c_int32), ("b_major", C.c_int32),
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("split_k", C.c_int32),
        ("epi", GemmEpilogue),
    ]


class CastItem(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("n", C.c_int64), ("src_dtype", C.c_int32), ("dst_dtype", C.c_int32)]


MAX_CAST_GROUP = 16


class AttnParams(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p),
        ("ldq", C.c_int64), ("ldk", C.c_int64), ("ldv", C.c_int64),
        ("mask", C.c_void_p), ("out", C.c_void_p), ("ldo", C.c_int64),
        ("lse", C.c_void_p), ("probs", C.c_void_p),
        ("B", C.c_int32), ("H", C.c_int32), ("Lq", C.c_int32), ("Lk", C.c_int32), ("dh", C.c_int32),
        ("scale", C.c_float), ("causal", C.c_int32),
        ("drop_p", C.c_float), ("drop_seed", C.c_uint64), ("drop_stream", C.c_uint64),
        ("dtype", C.c_int32),
        ("dout", C.c_void_p), ("dq", C.c_void_p), ("dk", C.c_void_p), ("dv", C.c_void_p),
        ("lddq", C.c_int64), ("lddk", C.c_int64), ("lddv", C.c_int64),
        ("delta", C.c_void_p), ("drop_mask", C.c_void_p), ("drop_mask_ready", C.c_int32),
    ]


class DropBitsItem(C.Structure):
    _fields_ = [("out", C.c_void_p), ("rows", C.c_int64), ("ld", C.c_int64), ("n", C.c_int32), ("kind", C.c_int32),
                ("p", C.c_float), ("seed", C.c_uint64), ("site", C.c_uint64)]


MAX_DROPBITS = 12


class SkinnyParams(C.Structure):
    _fields_ = [("x", C.c_void_p), ("ldx", C.c_int64), ("w", C.c_void_p), ("ldw", C.c_int64), ("bias", C.c_void_p),
                ("residual", C.c_void_p), ("ldr", C.c_int64), ("out", C.c_void_p * 3), ("ldo", C.c_int64 * 3),
                ("seg_cols", C.c_int32), ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("ksplit", C.c_int32),
                ("act", C.c_int32), ("ln_gamma", C.c_void_p), ("ln_beta", C.c_void_p), ("ln_eps", C.c_float),
                ("ws", C.c_void_p), ("counters", C.c_void_p)]


# name -> (restype, argtypes); must list every symbol include/m3ae_b200.h declares (tests check this).
SIGNATURES = {
    "m3ae_version": (C.c_int, []),
    "m3ae_last_error": (C.c_char_p, []),
    "m3ae_launch_count": (C.c_int64, []),
    "m3ae_gemm": (C.c_int, [C.POINTER(GemmProblem), C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "m3ae_attn_fwd": (C.c_int, [C.POINTER(AttnParams), C.c_void_p]),
    "m3ae_attn_bwd": (C.c_int, [C.POINTER(AttnParams), C.c_void_p]),
    "m3ae_attn_probs": (C.c_int, [C.POINTER(AttnParams), C.c_void_p]),
    "m3ae_activation_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "m3ae_activation_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "m3ae_bce_with_logits": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                                       C.c_void_p]),
    "m3ae_cross_entropy": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                     C.c_int64, C.c_int32, C.c_int64, C.c_void_p]),
    "m3ae_dropout_bits": (C.c_int, [C.POINTER(DropBitsItem), C.c_int32, C.c_void_p]),
    "m3ae_layernorm_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int64, C.c_int32, C.c_float, C.c_int32, C.c_void_p]),
    "m3ae_layernorm_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                     C.c_float, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]),
    "m3ae_colsum": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_int32, C.c_void_p]),
    "m3ae_cast": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]),
    "m3ae_cast_group": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p]),
    "m3ae_add_rowvec": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                  C.c_void_p]),
    "m3ae_dropout_apply": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_float, C.c_uint64,
                                     C.c_uint64, C.c_int32, C.c_void_p]),
    "m3ae_decode_attn": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                   C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                   C.c_float, C.c_int32, C.c_void_p]),
    "m3ae_argmax_rows": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "m3ae_decoder_embed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                     C.c_int32, C.c_void_p]),
    "m3ae_set_attention_backend": (C.c_int, [C.c_int32]),
    "m3ae_set_gemm_epilogue": (C.c_int, [C.c_int32]),
    "m3ae_argmax_partials": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "m3ae_set_gemm_skinny": (C.c_int, [C.c_int32]),
    "m3ae_set_gemm_reserved_sms": (C.c_int, [C.c_int32]),
    "m3ae_skinny_linear": (C.c_int, [C.POINTER(SkinnyParams), C.c_void_p]),
    "m3ae_skinny_linear_static": (C.c_int, [C.POINTER(SkinnyParams), C.c_void_p]),
    "m3ae_decoder_next_token": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "m3ae_set_pdl": (C.c_int, [C.c_int32]),
    "m3ae_set_gemm_cta_pair": (C.c_int, [C.c_int32]),
    "m3ae_set_layernorm_bwd_staging": (C.c_int, [C.c_int32]),
}

_lib = None


def load():
    """Load (once) and return the ctypes handle; raises if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"(or `make -C {os.path.join(_HERE, 'csrc')}`). There is no CPU/PyTorch fallback for the M3AE fusion path."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().m3ae_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")
