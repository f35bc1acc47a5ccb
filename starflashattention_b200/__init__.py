"""B200-native (sm_100a) FlashAttention forward and split-KV decode behind the StarFlashAttention interface.

Importing the package loads ``lib/libStarFlashAttention.so`` through its C ABI; there is no CPU or
PyTorch fallback, so a missing library is an ImportError (build with ``python starflashattention_b200/build.py``).
"""
from .interface import (DecoderStepGraph, compute_rotary_table, decode, decode_host, decode_num_splits, decode_paged, decoder_step_paged, fwd, fwd_host, fwd_host_last_pieces, fwd_host_plan, fwd_watchdog,
                        init_half_array, launch_count, mha_fwd_cuda, tuning_set)
from .multi_gpu import partition, shard_batch_heads

__all__ = ["fwd", "decode", "decode_paged", "decoder_step_paged", "fwd_host", "fwd_host_last_pieces", "fwd_host_plan", "decode_host", "mha_fwd_cuda", "compute_rotary_table", "init_half_array", "fwd_watchdog",
           "launch_count", "decode_num_splits", "tuning_set", "DecoderStepGraph", "partition", "shard_batch_heads"]
