"""Time the quantisation pre-pass operators one by one (development tool).
    python tools/bench_prepass.py [B H N D]
Inputs rotate through enough buffers to exceed the 126 MB L2, so every call reads HBM."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sageattention_b200 as S

B, H, N, D = (int(x) for x in sys.argv[1:5]) if len(sys.argv) >= 5 else (4, 32, 4096, 128)
dt = torch.float16
nbuf = max(2, int(300e6 // (B * H * N * D * 2)) + 1)
qs = [torch.randn(B, H, N, D, device="cuda", dtype=dt) for _ in range(nbuf)]
ks = [torch.randn(B, H, N, D, device="cuda", dtype=dt) + 1.0 for _ in range(nbuf)]
km = ks[0].mean(dim=2)
mb = B * H * N * D * 2 / 1e6


def timeit(name, fn, bytes_mb, reps=20):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    evs = []
    for i in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(i); b.record(); evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    t = ts[len(ts) // 2] * 1e3
    print(f"{name:44s} {t:8.1f} us   {bytes_mb / t * 1e3:7.0f} GB/s algorithmic")


print(f"B={B} H={H} N={N} D={D} fp16, {mb:.0f} MB per tensor, {nbuf} rotating buffers")
timeit("seq_mean(k)", lambda i: S.seq_mean(ks[i % nbuf]), mb)
timeit("per_thread_int8(q, k, km)  [Q + K]", lambda i: S.per_thread_int8(qs[i % nbuf], ks[i % nbuf], km), 2 * mb * 1.5)
timeit("per_warp_int8(q, k, km)    [Q + K]", lambda i: S.per_warp_int8(qs[i % nbuf], ks[i % nbuf], km), 2 * mb * 1.5)
timeit("smooth_quant_k_int8(k) per_thread [fused K]", lambda i: S.smooth_quant_k_int8(ks[i % nbuf], "per_thread"), mb * 1.5)
timeit("smooth_quant_k_int8(k) per_warp   [fused K]", lambda i: S.smooth_quant_k_int8(ks[i % nbuf], "per_warp"), mb * 1.5)
timeit("quant_q_int8(q) per_thread", lambda i: S.quant_q_int8(qs[i % nbuf], "per_thread"), mb * 1.5)
timeit("prepass_qk_int8(q, k) per_thread [one launch]", lambda i: S.prepass_qk_int8(qs[i % nbuf], ks[i % nbuf], "per_thread"), 2 * mb * 1.5)
timeit("prepass_qk_int8(q, k) per_warp   [one launch]", lambda i: S.prepass_qk_int8(qs[i % nbuf], ks[i % nbuf], "per_warp"), 2 * mb * 1.5)
timeit("prepass_qkv_fp8(q, k, v) per_thread [one launch]",
       lambda i: S.prepass_qkv_fp8(qs[i % nbuf], ks[i % nbuf], qs[(i + 1) % nbuf], "per_thread", _n_stat=(N + 127) // 128 * 128), 3 * mb * 1.5)
timeit("per_channel_fp8(v, smooth_v=False) [one launch]", lambda i: S.per_channel_fp8(qs[i % nbuf], smooth_v=False), mb * 1.5)
timeit("per_channel_fp8(v, smooth_v)", lambda i: S.per_channel_fp8(qs[i % nbuf], smooth_v=True), mb * 1.5)
timeit("torch copy (read+write)", lambda i: qs[(i + 1) % nbuf].copy_(qs[i % nbuf]), 2 * mb)
