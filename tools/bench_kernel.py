"""Kernel-only attention throughput sweep on one B200 (development tool; bench.py is the contract).

Protocol = the reference's own kernel benches (/root/reference/bench/bench_qk_int8_pv_fp8_cuda.py:72-98,
bench_qk_int8_pv_fp16_cuda.py:42-64): random int8 Q/K in [-95,95), random scales, randn V,
quantisation excluded, TOPS = 4*B*H*D*N^2 / (2 if causal) / t.  Scales here are |randn|*0.01 so the
softmax sees realistic magnitudes (the reference uses raw randn scales, negative ones included).

    python tools/bench_kernel.py [--pv f8,f16] [--d 128,64] [--n 1024,...] [--causal 0,1]
                                 [--baselines] [--reps 30] [--out gpurun_out/kernel_sweep.jsonl]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def time_fn(fn, reps, warm=5, graph=True):
    """Median / best device time of one fn() in ms.  Launches are replayed from a CUDA graph (several per
    replay when they are short) so that the number is the GPU's, not the Python caller's: one operator
    launch costs the host 60-100 us, more than the kernel itself takes at N <= 2K."""
    # warm up for at least 20 ms of device time: a short kernel timed right after a long, power-capped
    # configuration otherwise reads low (clock recovery), as does the first configuration of a run
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    spent, n = 0.0, 0
    while n < warm or (spent < 20.0 and n < 2000):
        t0.record()
        fn()
        t1.record()
        torch.cuda.synchronize()
        spent += t0.elapsed_time(t1)
        n += 1
    inner, run = 1, fn
    if graph:
        try:
            inner = 8 if spent / n < 1.0 else 1
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(inner):
                    fn()
            run = g.replay
            run()
            torch.cuda.synchronize()
        except Exception as e:  # launches outside the capturing stream etc.: time the calls directly
            print("graph capture failed, timing direct calls:", type(e).__name__, file=sys.stderr)
            torch.cuda.synchronize()
            inner, run = 1, fn
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record()
        run()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) / inner for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pv", default="f8,f16")
    ap.add_argument("--d", default="128,64")
    ap.add_argument("--n", default="1024,2048,4096,8192,16384,32768")
    ap.add_argument("--causal", default="0,1")
    ap.add_argument("--gran", default="per_thread")
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--baselines", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "kernel_sweep.jsonl"))
    args = ap.parse_args()

    import sageattention_b200 as sb
    from sageattention_b200 import core
    import time
    import bench
    dev = "cuda:0"
    # every line carries the SM clocks / throttle reasons sampled while it was measured (numbers without a clock
    # record are not evidence)
    sampler = bench.ClockSampler(0)
    sampler.start()
    sampler.wait_first()
    B, H = args.batch, args.heads
    gran = 3 if args.gran == "per_thread" else 2
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    out = open(args.out, "a")

    t_mark = [time.monotonic()]

    def emit(rec):
        now = time.monotonic()
        rec["clocks"] = sampler.window(t_mark[0], now)
        t_mark[0] = now
        line = json.dumps(rec)
        print(line, flush=True)
        out.write(line + "\n")
        out.flush()

    ref80 = None
    if args.baselines:
        from oracle import build_ref
        try:
            ref80 = build_ref.load("sm80")
        except Exception as e:
            print("no sm80 ref:", e)
        try:
            from flash_attn import flash_attn_func
        except Exception as e:
            flash_attn_func = None
            print("no flash_attn:", e)

    for d in [int(x) for x in args.d.split(",")]:
        for causal in [bool(int(x)) for x in args.causal.split(",")]:
            for n in [int(x) for x in args.n.split(",")]:
                flops = 4 * H * B * d * n * n / (2 if causal else 1)
                q8 = torch.randint(-95, 95, (B, H, n, d), dtype=torch.int8, device=dev)
                k8 = torch.randint(-95, 95, (B, H, n, d), dtype=torch.int8, device=dev)
                nqs = (n + 127) // 128 * 4 * (8 if gran == 3 else 1)
                nks = (n + 63) // 64 * (4 if gran == 3 else 1)
                qs = torch.randn(B, H, nqs, device=dev).abs() * 0.01 + 0.005
                ks = torch.randn(B, H, nks, device=dev).abs() * 0.01 + 0.005
                v = torch.randn(B, H, n, d, dtype=torch.float16, device=dev)
                o = torch.empty(B, H, n, d, dtype=torch.float16, device=dev)
                sm = d ** -0.5
                for pv in args.pv.split(","):
                    if pv == "f8":
                        npad = (n + 127) // 128 * 128
                        vq = torch.randn(B, H, d, npad, device=dev).to(torch.float8_e4m3fn)
                        vs = torch.randn(B, H, d, device=dev).abs()
                    else:
                        vq, vs = v, None
                    fn = lambda: core._launch(pv == "f8", q8, k8, vq, o, qs, ks, vs, None, "HND", causal, gran, 32,
                                              sm, False)
                    med, best = time_fn(fn, args.reps)
                    emit(dict(impl="sage_b200", pv=pv, d=d, causal=causal, n=n, ms=med, ms_best=best,
                              tops=flops / med * 1e-9, tops_best=flops / best * 1e-9))
                if args.baselines:
                    if ref80 is not None and n <= 16384:
                        fn = lambda: ref80.qk_int8_sv_f16_accum_f32_attn(q8, k8, v, o, qs, ks, 1, int(causal), gran, sm, 0)
                        med, best = time_fn(fn, max(5, args.reps // 3), graph=False)
                        emit(dict(impl="ref_sm80_f32", d=d, causal=causal, n=n, ms=med, tops=flops / med * 1e-9))
                        fn = lambda: ref80.qk_int8_sv_f16_accum_f16_attn(q8, k8, v, o, qs, ks, 1, int(causal), gran, sm, 0)
                        med, best = time_fn(fn, max(5, args.reps // 3), graph=False)
                        emit(dict(impl="ref_sm80_f16", d=d, causal=causal, n=n, ms=med, tops=flops / med * 1e-9))
                    qb = torch.randn(B, n, H, d, dtype=torch.bfloat16, device=dev)
                    kb = torch.randn(B, n, H, d, dtype=torch.bfloat16, device=dev)
                    vb = torch.randn(B, n, H, d, dtype=torch.bfloat16, device=dev)
                    if flash_attn_func is not None:
                        fn = lambda: flash_attn_func(qb, kb, vb, causal=causal)
                        med, best = time_fn(fn, max(5, args.reps // 3))
                        emit(dict(impl="flash_attn2_bf16", d=d, causal=causal, n=n, ms=med, tops=flops / med * 1e-9))
                    qh, kh, vh = (t.transpose(1, 2) for t in (qb, kb, vb))
                    try:
                        fn = lambda: torch.nn.functional.scaled_dot_product_attention(qh, kh, vh, is_causal=causal)
                        med, best = time_fn(fn, max(5, args.reps // 3))
                        emit(dict(impl="torch_sdpa_bf16", d=d, causal=causal, n=n, ms=med, tops=flops / med * 1e-9))
                    except Exception as e:
                        print("sdpa failed", e)
                    del qb, kb, vb


if __name__ == "__main__":
    main()
