"""Host-side cost of one operator call (development tool): time N back-to-back calls without
synchronising (the CPU only queues work) against the GPU time of the same calls."""
import os, sys, time, cProfile, pstats
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sageattention_b200 as S

q, k, v = (torch.randn(4, 32, 4096, 128, device="cuda", dtype=torch.float16) for _ in range(3))
fn = getattr(S, sys.argv[1]) if len(sys.argv) > 1 else S.sageattn_qk_int8_pv_fp16_cuda
print("entry point:", fn.__name__)
for _ in range(5):
    fn(q, k, v)
torch.cuda.synchronize()
n = 100
# small problem: GPU time negligible, so wall time = host time
qs, ks, vs = (torch.randn(1, 2, 256, 128, device="cuda", dtype=torch.float16) for _ in range(3))
for _ in range(5):
    fn(qs, ks, vs)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(n):
    fn(qs, ks, vs)
t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"host time per call (tiny problem, no sync): {(t1 - t0) / n * 1e6:.1f} us")
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    fn(qs, ks, vs)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(int(sys.argv[2]) if len(sys.argv) > 2 else 18)
