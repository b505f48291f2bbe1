import sys, torch
sys.path.insert(0, "/root/repo")
import sageattention_b200 as S
def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3
for H in (5, 40):
    for N in (75600, 119056 if H == 5 else 75600):
        q, k, v = (torch.randn(1, H, N, 128, device="cuda", dtype=torch.bfloat16) for _ in range(3))
        nst = (N + 127) // 128 * 128
        c_qkv = t(lambda: S.prepass_qkv_fp8(q, k, v, "per_thread", _n_stat=nst))
        c_qk = t(lambda: S.prepass_qk_int8(q, k, "per_thread"))
        f_mean = t(lambda: S.seq_mean(k))
        km = S.seq_mean(k)
        f_qk = t(lambda: S.per_thread_int8(q, k, km))
        c_v = t(lambda: S.per_channel_fp8(v, smooth_v=False, _n_stat=nst))
        f_v = t(lambda: S.per_channel_fp8(v, smooth_v=True, _n_stat=nst))   # flat grid (exact-order stats + flat conversion)
        print(f"H={H} N={N}: cluster qkv {c_qkv:7.1f} us | cluster qk {c_qk:7.1f} | flat mean {f_mean:6.1f} + flat q,k quant {f_qk:6.1f} = {f_mean+f_qk:7.1f} | V cluster {c_v:7.1f}  V flat(smooth_v kernels) {f_v:7.1f}", flush=True)
        del q, k, v
