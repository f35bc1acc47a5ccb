"""Timing of the drop-in decoder step (mha_fwd_cuda: RoPE + append + split-KV attention) on a long cache:
fused single kernel vs the three-launch form, eager vs CUDA-graph replay.

    python tools/decoder_step_time.py [batch] [heads] [cache_len] [layers]
"""
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import starflashattention_b200 as sfa  # noqa: E402


def timed(fn, iters, flush):
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts), min(ts)


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    H = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    M = int(sys.argv[3]) if len(sys.argv) > 3 else 32768
    L = int(sys.argv[4]) if len(sys.argv) > 4 else 4
    D, dt = 128, torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(0)
    kc = torch.randn((B, L, M, H, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
    vc = torch.randn((B, L, M, H, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
    qkv = torch.randn((L, B, 3, H, D), generator=g, device="cuda", dtype=torch.float32).to(dt)
    cos, sin = sfa.compute_rotary_table(M, D, dtype=dt)
    seq = torch.full((B,), M - 64, dtype=torch.int32, device="cuda")
    o = torch.empty((L, B, H, D), dtype=dt, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    bytes_step = 2.0 * B * L * (M - 63) * H * D * 2   # K + V read once per layer

    def eager():
        for layer in range(L):
            sfa.mha_fwd_cuda(qkv[layer], None, None, None, kc, vc, seq, o[layer], B, M, H, D, D, M, L, layer,
                             rotary_cos_table=cos, rotary_sin_table=sin)

    for fused in ("1", "0"):
        sfa.tuning_set("decoder_fused", int(fused))
        for _ in range(3):
            eager()
        n0 = sfa.launch_count()
        med, mn = timed(eager, 20, flush)
        per = (sfa.launch_count() - n0) // 20
        print(f"[decoder_step] eager fused={fused}: median {med:.1f} us  min {mn:.1f} us  {bytes_step / med / 1e3:.0f} GB/s  "
              f"({per} launches per step, {L} layers)", flush=True)
    sfa.tuning_set("decoder_fused", 1)
    gr = sfa.DecoderStepGraph(kc, vc, seq, rotary_cos_table=cos, rotary_sin_table=sin)
    gr.qkv.copy_(qkv)

    def replay():
        gr.graph.replay()
        gr.seq_len.copy_(seq)      # keep the position fixed so every replay does the same work (outside the graph)

    for _ in range(3):
        replay()
    med, mn = timed(lambda: gr.graph.replay(), 1, flush)
    gr.seq_len.copy_(seq)
    ts = []
    for _ in range(20):
        flush.zero_(); gr.seq_len.copy_(seq)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.graph.replay(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    med = statistics.median(ts)
    print(f"[decoder_step] graph replay fused=1: median {med:.1f} us  min {min(ts):.1f} us  {bytes_step / med / 1e3:.0f} GB/s  "
          f"(1 graph launch per step, {L} layers)", flush=True)
    # small-cache latency: where launch overhead is what matters
    seq.fill_(255)
    gr2 = sfa.DecoderStepGraph(kc, vc, seq, rotary_cos_table=cos, rotary_sin_table=sin)
    for name, fn in (("eager", eager), ("graph", lambda: (gr2.seq_len.copy_(seq), gr2.graph.replay()))):
        for _ in range(3):
            fn()
        med, mn = timed(fn, 50, None)
        print(f"[decoder_step] short cache (256 keys) {name}: median {med:.1f} us  min {mn:.1f} us per {L}-layer step", flush=True)


if __name__ == "__main__":
    main()
