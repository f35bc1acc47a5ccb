"""ctypes binding of libStarFlashAttention.so (the C ABI declared in include/star_flash_attn.h).

There is no CPU fallback: if the shared library has not been built, importing this module raises.
Build it with ``python starflashattention_b200/build.py`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libStarFlashAttention.so")
if os.environ.get("SFA_PROBES_LIB") == "1":      # tuning tools only: the -DSFA_PROBES build (build.py --probes)
    LIB_PATH = os.path.join(_PKG, "lib", "libStarFlashAttention_probes.so")

SFA_FP16, SFA_BF16 = 0, 1
SFA_OK, SFA_ERR_INVALID, SFA_ERR_UNSUPPORTED, SFA_ERR_CUDA, SFA_ERR_NO_DEVICE = 0, -1, -2, -3, -4

# Every symbol include/star_flash_attn.h declares; tests check the library exports all of them.
EXPORTED_SYMBOLS = (
    "sfa_fwd", "sfa_decode", "sfa_decode_static", "sfa_decode_static_scratch_bytes", "sfa_decode_paged", "sfa_fwd_host", "sfa_decode_host", "sfa_fwd_host_plan", "sfa_fwd_host_last_pieces", "sfa_fwd_host_trace", "sfa_copy_2d", "sfa_decode_num_splits", "sfa_decode_workspace_bytes", "sfa_decoder_step", "sfa_decoder_step_workspace_bytes", "sfa_decoder_step_paged",
    "sfa_rotary_table", "sfa_fill", "sfa_last_error", "sfa_abi_version", "sfa_launch_count",
    "sfa_fwd_watchdog", "sfa_fwd_probe_read", "sfa_release_workspaces", "sfa_tuning_set",
    "sfa_decode_sharded", "sfa_decode_sharded_emulated", "sfa_shard_merge", "sfa_shard_workspace_bytes", "sfa_peer_alloc", "sfa_peer_open",
    "sfa_peer_close", "sfa_peer_free", "sfa_shard_watchdog",
)
# C++-mangled drop-ins of the reference's src/flash_attn.h:7-11.
MANGLED_DROP_INS = (
    "_Z17run_flash_decoderI6__halfEvR19Flash_decoder_inputR20Flash_decoder_paramsP11CUstream_st",
    "_Z20compute_rotary_tableI6__halfEvPT_S2_ii",
    "_Z15init_half_arrayP6__halfS_iii",
)


class FwdArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("o", C.c_void_p), ("lse", C.c_void_p),
        ("batch", C.c_int32), ("seqlen_q", C.c_int32), ("seqlen_k", C.c_int32),
        ("heads_q", C.c_int32), ("heads_kv", C.c_int32), ("head_dim", C.c_int32),
        ("q_batch_stride", C.c_int64), ("q_row_stride", C.c_int64), ("q_head_stride", C.c_int64),
        ("k_batch_stride", C.c_int64), ("k_row_stride", C.c_int64), ("k_head_stride", C.c_int64),
        ("v_batch_stride", C.c_int64), ("v_row_stride", C.c_int64), ("v_head_stride", C.c_int64),
        ("o_batch_stride", C.c_int64), ("o_row_stride", C.c_int64), ("o_head_stride", C.c_int64),
        ("softmax_scale", C.c_float), ("causal", C.c_int32), ("dtype", C.c_int32),
    ]


class DecodeArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p), ("k_cache", C.c_void_p), ("v_cache", C.c_void_p), ("seqlens", C.c_void_p),
        ("o", C.c_void_p), ("lse", C.c_void_p),
        ("batch", C.c_int32), ("seqlen_k", C.c_int32), ("heads_q", C.c_int32), ("heads_kv", C.c_int32),
        ("head_dim", C.c_int32),
        ("q_batch_stride", C.c_int64), ("q_head_stride", C.c_int64),
        ("k_batch_stride", C.c_int64), ("k_row_stride", C.c_int64), ("k_head_stride", C.c_int64),
        ("v_batch_stride", C.c_int64), ("v_row_stride", C.c_int64), ("v_head_stride", C.c_int64),
        ("o_batch_stride", C.c_int64), ("o_head_stride", C.c_int64),
        ("softmax_scale", C.c_float), ("num_splits", C.c_int32), ("dtype", C.c_int32),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
    ]


class FwdHostArgs(C.Structure):
    _fields_ = [("t", FwdArgs), ("pieces", C.c_int32)]


class DecodeHostArgs(C.Structure):
    _fields_ = [("t", DecodeArgs), ("pieces", C.c_int32)]


class DecodePagedArgs(C.Structure):
    _fields_ = [
        ("base", DecodeArgs), ("block_table", C.c_void_p), ("block_table_stride", C.c_int64),
        ("page_size", C.c_int32), ("num_pages", C.c_int32),
    ]


class DecoderStepPagedArgs(C.Structure):
    _fields_ = [
        ("qkv", C.c_void_p), ("o", C.c_void_p), ("lse", C.c_void_p), ("seq_len", C.c_void_p),
        ("k_pool", C.c_void_p), ("v_pool", C.c_void_p), ("block_table", C.c_void_p),
        ("rotary_cos_table", C.c_void_p), ("rotary_sin_table", C.c_void_p),
        ("page_stride", C.c_int64), ("row_stride", C.c_int64), ("head_stride", C.c_int64),
        ("block_table_stride", C.c_int64), ("qkv_stride", C.c_int64),
        ("batch_size", C.c_int32), ("max_len", C.c_int32), ("heads_q", C.c_int32), ("heads_kv", C.c_int32),
        ("head_dim", C.c_int32), ("page_size", C.c_int32), ("num_pages", C.c_int32),
        ("rotary_embedding_dim", C.c_int32), ("softmax_scale", C.c_float), ("num_splits", C.c_int32),
        ("dtype", C.c_int32),
    ]


class DecoderStepArgs(C.Structure):
    _fields_ = [
        ("qkv", C.c_void_p), ("o", C.c_void_p), ("seq_len", C.c_void_p),
        ("k_cache_table", C.c_void_p), ("v_cache_table", C.c_void_p),
        ("rotary_cos_table", C.c_void_p), ("rotary_sin_table", C.c_void_p), ("lse", C.c_void_p),
        ("batch_size", C.c_int32), ("memory_max_len", C.c_int32), ("num_heads", C.c_int32), ("head_dim", C.c_int32),
        ("head_dim_inv", C.c_float),
        ("rotary_embedding_dim", C.c_int32), ("qkv_stride", C.c_int32), ("num_layer", C.c_int32),
        ("idx_layer", C.c_int32), ("num_splits", C.c_int32), ("dtype", C.c_int32),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
    ]


class ShardComm(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32), ("step", C.c_uint32), ("mode", C.c_int32),
        ("peer_ws", C.c_void_p * 8), ("payload", C.c_void_p),
    ]


class StarFlashAttentionError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libStarFlashAttention error {code}: {message}")
        self.code = code


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing. Build it with `python starflashattention_b200/build.py`; "
            "this package has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.sfa_fwd.argtypes = [C.POINTER(FwdArgs), C.c_void_p]
    lib.sfa_fwd.restype = C.c_int
    lib.sfa_decode.argtypes = [C.POINTER(DecodeArgs), C.c_void_p]
    lib.sfa_decode.restype = C.c_int
    lib.sfa_decode_static.argtypes = [C.POINTER(DecodeArgs), C.c_void_p]
    lib.sfa_decode_static.restype = C.c_int
    lib.sfa_decode_static_scratch_bytes.argtypes = []
    lib.sfa_decode_static_scratch_bytes.restype = C.c_size_t
    lib.sfa_decode_paged.argtypes = [C.POINTER(DecodePagedArgs), C.c_void_p]
    lib.sfa_decode_paged.restype = C.c_int
    lib.sfa_fwd_host.argtypes = [C.POINTER(FwdHostArgs), C.c_void_p]
    lib.sfa_fwd_host.restype = C.c_int
    lib.sfa_decode_host.argtypes = [C.POINTER(DecodeHostArgs), C.c_void_p]
    lib.sfa_decode_host.restype = C.c_int
    lib.sfa_fwd_host_plan.argtypes = [C.POINTER(FwdHostArgs), C.POINTER(C.c_int32), C.c_int32]
    lib.sfa_fwd_host_plan.restype = C.c_int32
    lib.sfa_fwd_host_last_pieces.argtypes = []
    lib.sfa_fwd_host_last_pieces.restype = C.c_int32
    lib.sfa_fwd_host_trace.argtypes = [C.POINTER(C.c_float), C.c_int32]
    lib.sfa_fwd_host_trace.restype = C.c_int32
    lib.sfa_copy_2d.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int32, C.c_void_p]
    lib.sfa_copy_2d.restype = C.c_int
    lib.sfa_decode_num_splits.argtypes = [C.POINTER(DecodeArgs)]
    lib.sfa_decode_num_splits.restype = C.c_int
    lib.sfa_decode_workspace_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    lib.sfa_decode_workspace_bytes.restype = C.c_size_t
    lib.sfa_decoder_step.argtypes = [C.POINTER(DecoderStepArgs), C.c_void_p]
    lib.sfa_decoder_step.restype = C.c_int
    lib.sfa_decoder_step_workspace_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32]
    lib.sfa_decoder_step_workspace_bytes.restype = C.c_size_t
    lib.sfa_decoder_step_paged.argtypes = [C.POINTER(DecoderStepPagedArgs), C.c_void_p]
    lib.sfa_decoder_step_paged.restype = C.c_int
    lib.sfa_rotary_table.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    lib.sfa_rotary_table.restype = C.c_int
    lib.sfa_fill.argtypes = [C.c_void_p, C.c_float, C.c_int64, C.c_int32, C.c_void_p]
    lib.sfa_fill.restype = C.c_int
    lib.sfa_last_error.restype = C.c_char_p
    lib.sfa_abi_version.restype = C.c_int
    lib.sfa_launch_count.restype = C.c_uint64
    lib.sfa_fwd_watchdog.argtypes = [C.c_int]
    lib.sfa_fwd_watchdog.restype = C.c_int
    lib.sfa_fwd_probe_read.argtypes = [C.POINTER(C.c_uint64), C.c_int32]
    lib.sfa_fwd_probe_read.restype = C.c_int
    lib.sfa_release_workspaces.restype = None
    lib.sfa_tuning_set.argtypes = [C.c_char_p, C.c_int32]
    lib.sfa_tuning_set.restype = C.c_int
    lib.sfa_decode_sharded.argtypes = [C.POINTER(DecodeArgs), C.POINTER(ShardComm), C.c_void_p]
    lib.sfa_decode_sharded.restype = C.c_int
    lib.sfa_decode_sharded_emulated.argtypes = [C.POINTER(DecodeArgs), C.c_int32, C.POINTER(C.c_void_p), C.c_uint32, C.c_void_p]
    lib.sfa_decode_sharded_emulated.restype = C.c_int
    lib.sfa_shard_merge.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                    C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
    lib.sfa_shard_merge.restype = C.c_int
    lib.sfa_shard_workspace_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    lib.sfa_shard_workspace_bytes.restype = C.c_size_t
    lib.sfa_peer_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]
    lib.sfa_peer_alloc.restype = C.c_int
    lib.sfa_peer_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.sfa_peer_open.restype = C.c_int
    lib.sfa_peer_close.argtypes = [C.c_void_p]
    lib.sfa_peer_close.restype = C.c_int
    lib.sfa_peer_free.argtypes = [C.c_void_p]
    lib.sfa_peer_free.restype = C.c_int
    lib.sfa_shard_watchdog.argtypes = [C.c_int]
    lib.sfa_shard_watchdog.restype = C.c_int
    return lib


lib = _load()


def check(rc: int) -> None:
    if rc != 0:
        raise StarFlashAttentionError(rc, (lib.sfa_last_error() or b"").decode())
