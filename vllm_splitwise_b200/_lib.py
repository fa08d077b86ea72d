"""ctypes binding of libb200kv.so (the C ABI declared in include/b200_kv.h).

This is the only place that touches the shared library.  It fails loudly when the library is
missing -- there is no fallback implementation of any op.
"""
import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_float, c_int32, c_int64, c_ubyte, c_uint64,
                    c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200_KV_LIB points at another build of the same library (development: `make FAST=1 OUTDIR=...`)
LIB_PATH = os.environ.get("B200_KV_LIB") or os.path.join(_HERE, "lib", "libb200kv.so")

B200_OK = 0
ERR_HEAD_SIZE, ERR_BLOCK_SIZE, ERR_DTYPE, ERR_ARG, ERR_ALIGN, ERR_CUDA, ERR_DEVICE, ERR_TIMEOUT = (
    -1, -2, -3, -4, -5, -6, -7, -8)
F32, F16, BF16, U8_E5M2 = 0, 1, 2, 3
SWAP_D2D, SWAP_D2H, SWAP_H2D = 0, 1, 2
IPC_HANDLE_BYTES = 64


class PagedAttnArgs(Structure):
    """Mirror of b200_paged_attn_args_t (field order and types must match the header)."""
    _fields_ = [
        ("out", c_void_p),
        ("query", c_void_p),
        ("key_cache", c_void_p),
        ("value_cache", c_void_p),
        ("block_tables", c_void_p),
        ("context_lens", c_void_p),
        ("alibi_slopes", c_void_p),
        ("scale", c_float),
        ("num_seqs", c_int32),
        ("num_heads", c_int32),
        ("num_kv_heads", c_int32),
        ("head_size", c_int32),
        ("block_size", c_int32),
        ("max_context_len", c_int32),
        ("max_num_blocks_per_seq", c_int32),
        ("q_stride", c_int64),
        ("kv_block_stride", c_int64),
        ("kv_head_stride", c_int64),
        ("query_dtype", c_int32),
        ("cache_dtype", c_int32),
        ("exp_sums", c_void_p),
        ("max_logits", c_void_p),
        ("tmp_out", c_void_p),
        ("max_num_partitions", c_int32),
        ("workspace", c_void_p),
        ("workspace_bytes", c_int64),
        ("device", c_int32),
        ("reserved", c_int32),
    ]


class ContextAttnArgs(Structure):
    """Mirror of b200_context_attn_args_t (field order and types must match the header)."""
    _fields_ = ([(n, c_void_p) for n in ("out", "query", "key", "value", "key_cache", "value_cache", "block_tables",
                                         "start_loc", "seq_lens", "ctx_lens", "alibi_slopes")]
                + [("scale", c_float)]
                + [(n, c_int32) for n in ("batch", "num_heads", "num_kv_heads", "head_size", "block_size",
                                          "max_input_len", "block_table_stride")]
                + [(n, c_int64) for n in ("q_stride_t", "q_stride_h", "k_stride_t", "k_stride_h", "v_stride_t",
                                          "v_stride_h", "o_stride_t", "o_stride_h", "kc_block_stride",
                                          "kc_head_stride", "vc_block_stride", "vc_head_stride")]
                + [("dtype", c_int32), ("device", c_int32)])


class DecodeStepArgs(Structure):
    """Mirror of b200_decode_step_args_t (field order and types must match the header)."""
    _fields_ = [
        ("attn", PagedAttnArgs),
        ("state_block_tables", c_void_p),
        ("state_seq_lens", c_void_p),
        ("batch_rows", c_void_p),
        ("state_max_blocks_per_seq", c_int32),
        ("sliding_window", c_int32),
        ("positions", c_void_p),
        ("slot_mapping", c_void_p),
        ("key", c_void_p),
        ("value", c_void_p),
        ("cos_sin_cache", c_void_p),
        ("rot_dim", c_int32),
        ("is_neox", c_int32),
        ("key_stride", c_int64),
        ("value_stride", c_int64),
        ("num_blocks", c_int64),
    ]


# name -> (restype, argtypes); every symbol include/b200_kv.h declares
SIGNATURES = {
    "b200_kv_strerror": (c_char_p, [c_int32]),
    "b200_kv_last_cuda_error": (c_char_p, []),
    "b200_kv_abi_version": (c_int32, []),
    "b200_kv_launch_count": (c_uint64, []),
    "b200_paged_attention_v1": (c_int32, [POINTER(PagedAttnArgs), c_void_p]),
    "b200_paged_attention_v2": (c_int32, [POINTER(PagedAttnArgs), c_void_p]),
    "b200_paged_attention_workspace_bytes": (c_int64, [POINTER(PagedAttnArgs)]),
    "b200_paged_attention_v2_static_scratch": (c_int32, [POINTER(PagedAttnArgs), c_void_p]),
    "b200_paged_attention_v2_static_scratch_bytes": (c_int64, [POINTER(PagedAttnArgs)]),
    "b200_context_attention_fwd": (c_int32, [POINTER(ContextAttnArgs), c_void_p]),
    "b200_allreduce_oneshot": (c_int32, [POINTER(c_int64), POINTER(c_int64), POINTER(c_int64), c_void_p, c_void_p,
                                         c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_int32, c_uint64,
                                         c_int64, c_int32, c_void_p]),
    "b200_allreduce_twostage": (c_int32, [POINTER(c_int64), POINTER(c_int64), POINTER(c_int64), POINTER(c_int64),
                                          POINTER(c_int64), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_int64, c_int32, c_int32, c_int32, c_uint64, c_int64, c_void_p]),
    "b200_block_tables_append": (c_int32, [c_void_p, c_void_p, c_int32, c_int32, c_void_p]),
    "b200_decode_metadata_step": (c_int32, [c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_void_p,
                                            c_void_p, c_void_p, c_void_p, c_int32, c_void_p]),
    "b200_decode_step": (c_int32, [POINTER(DecodeStepArgs), c_void_p]),
    "b200_reshape_and_cache": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32,
                                         c_int32, c_int32, c_int32, c_int32, c_int64, c_int64,
                                         c_int64, c_int32, c_int32, c_void_p]),
    "b200_copy_blocks": (c_int32, [POINTER(c_int64), POINTER(c_int64), c_int32, POINTER(c_int64),
                                   c_int32, c_int64, c_void_p]),
    "b200_swap_blocks": (c_int32, [c_void_p, c_void_p, c_int64, POINTER(c_int64), c_int32, c_int32,
                                   c_int32, c_int32, c_void_p]),
    "b200_convert_fp8_e5m2": (c_int32, [c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p]),
    "b200_rotary_embedding": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32,
                                        c_int32, c_int32, c_int64, c_int64, c_int32, c_int32, c_void_p]),
    "b200_rotary_embedding_reshape_and_cache": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                                          c_int32, c_int32, c_int32, c_int32, c_int64, c_int64,
                                                          c_int64, c_int32, c_void_p, c_void_p, c_void_p, c_int32,
                                                          c_int32, c_int64, c_int32, c_int32, c_void_p]),
    "b200_get_device_attribute": (c_int32, [c_int32, c_int32]),
    "b200_get_max_shared_memory_per_block_device_attribute": (c_int32, [c_int32]),
    "b200_kv_ipc_export": (c_int32, [c_void_p, POINTER(c_ubyte), POINTER(c_int64),
                                     POINTER(c_int64)]),
    "b200_kv_ipc_open": (c_int32, [POINTER(c_ubyte), POINTER(c_void_p)]),
    "b200_kv_ipc_close": (c_int32, [c_void_p]),
    "b200_kv_enable_peer_access": (c_int32, [c_int32, c_int32]),
    "b200_kv_put_blocks": (c_int32, [c_void_p, c_void_p, c_int32, POINTER(c_int32), c_int32,
                                     c_int64, c_int32, c_void_p]),
    "b200_kv_put_configure": (c_int32, [c_int32, c_int32]),
    "b200_kv_put_config": (c_int32, [POINTER(c_int32), POINTER(c_int32)]),
    "b200_kv_put_table_build": (c_int32, [POINTER(c_int32), c_int32, c_int64, c_void_p, POINTER(c_int32),
                                          POINTER(c_int64)]),
    "b200_kv_put_blocks_table": (c_int32, [c_void_p, c_void_p, c_int32, c_void_p, c_int32, c_int32, c_int64, c_int64,
                                           c_int32, c_void_p]),
    "b200_kv_signal": (c_int32, [c_void_p, c_uint64, c_void_p]),
    "b200_kv_wait": (c_int32, [c_void_p, c_uint64, c_int64, c_void_p, c_void_p]),
}

_lib = None


def load():
    """Load libb200kv.so once; raise ImportError (never fall back) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (or `make -C vllm_splitwise_b200/csrc`). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status, what, detail=""):
    """Map a b200_status_t to the exception the reference op would raise (RuntimeError)."""
    if status == B200_OK:
        return
    lib = load()
    msg = lib.b200_kv_strerror(status).decode()
    if status == ERR_CUDA:
        msg += ": " + lib.b200_kv_last_cuda_error().decode()
    if detail:
        msg += ": " + str(detail)
    raise RuntimeError(f"{what}: {msg}")
