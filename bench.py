"""bench.py - the measurement contract for the SageAttention hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME] [--no-sweep] [--no-sp]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (quantisation pre-pass + fused INT8-QK / online-softmax / PV
attention kernel) over one batch of synthetic input.  Default workload = BASELINE.json configs[1]:
`sageattn_qk_int8_pv_fp16_cuda`, non-causal, B=4 H=32 N=4096 hd128 fp16 (the reference's
bench_qk_int8_pv_fp16_cuda shape).  Other named workloads: c3_fp8 (fp8 PV, same shape), hunyuan,
wan (DiT layer shapes; under torchrun they run head-sharded: independent heads per rank, no
collective - SURVEY 8e).

metric  = attention TOPS = 4*B*H*D*Nq*Nk / (2 if causal) / t       (reference formula,
          /root/reference/bench/bench_qk_int8_pv_fp16_cuda.py:42)
value   = whole-job TOPS of the full operator call with q/k/v already resident in HBM
e2e     = same metric through the public API with HOST (pinned) q/k/v: H2D of the inputs and D2H of
          the output are inside the timed region of every step
roofline= the attention kernel alone (CUDA events around its launch inside the timed steps)
sweep   = (N=1) the metric's whole range measured in the same run: hd {64,128} x {full, causal} x N 1K..32K x PV
          {fp8, f16}, kernel-only and whole-operator (public API, no graph replay) TOPS, clocks per entry
sp      = the Wan2.1 DiT layer (B1 H40 N75600 hd128 bf16): on one GPU at N=1, under Ulysses sequence parallelism
          (NCCL all-to-all) at N>1 with the collective timed alone and the efficiency against one GPU
roofline.measured_ceilings = dense INT8 / FP8 / bf16 GEMM throughput of this GPU measured in this run
cpu_baseline / --impl reference = the CPU restatement of the reference algorithm (oracle/, "port":
          the reference has no CPU implementation of this path) on the host cores, on a
          bounded sample of the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (api, B, H, N, D, causal, dtype)
    "c2_fp16": ("sageattn_qk_int8_pv_fp16_cuda", 4, 32, 4096, 128, False, "float16"),
    "c2_fp16_bf16in": ("sageattn_qk_int8_pv_fp16_cuda", 4, 32, 4096, 128, False, "bfloat16"),
    "c3_fp8": ("sageattn_qk_int8_pv_fp8_cuda", 4, 32, 4096, 128, False, "float16"),
    "c3_fp8_8k": ("sageattn_qk_int8_pv_fp8_cuda", 4, 32, 8192, 128, False, "bfloat16"),
    "hunyuan": ("sageattn", 1, 24, 119056, 128, False, "bfloat16"),
    "wan": ("sageattn", 1, 40, 75600, 128, False, "bfloat16"),
}
# dram__bytes_read.sum + dram__bytes_write.sum of the attention kernel, one `ncu --set full` capture
# per workload of the current (unified) kernel (profiles/r2n_ncu_attn_summary.txt, end of round 2); algorithmic bytes at
# C2 fp16 = 402.7e6
NCU_TRAFFIC = {"c2_fp16": 376.61e6, "c3_fp8": 305.46e6}
NCU_TRAFFIC_SOURCE = {"c2_fp16": "ncu capture profiles/r2n_ncu_attn_summary.txt, round 2 (not measured in this run)",
                      "c3_fp8": "ncu capture profiles/r2n_ncu_attn_summary.txt, round 2 (not measured in this run)"}
METRIC = "attention TOPS @hd64/hd128 seq 1K-32K (causal/non-causal), % of INT8/FP8 tensor peak"


def flops_of(b, h, n, d, causal):
    return 4.0 * b * h * d * n * n / (2 if causal else 1)


def peaks():
    """Tensor-roofline denominators.  MEASURED_PEAKS.json holds the driver's measured dense bf16
    cuBLAS throughput; tcgen05 runs kind::i8 / kind::f8f6f4 at twice the kind::f16 rate, so the
    time-weighted ceiling of a kernel with half its flops in INT8 QK^T and half in f16 (fp8) PV is
    bf16 * 4/3 (bf16 * 2)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return float(j["bf16_tflops"]), float(j.get("hbm_gbs", 6650.0)), "measured"
    return 1590.0, 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.monotonic(), line.strip()))

    def wait_first(self, timeout=5.0):
        """nvidia-smi needs a few hundred ms to come up: block until it delivers."""
        t0 = time.monotonic()
        while self.proc is not None and not self.lines and time.monotonic() - t0 < timeout:
            time.sleep(0.01)

    def window(self, t_begin=None, t_end=None):
        """Summary of the samples taken in [t_begin, t_end] (monotonic clock); the sampler keeps running."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # only samples taken inside the window count; a window shorter than the 50 ms sampling period falls
        # back to the samples within 100 ms of it
        lines = list(self.lines)
        if t_begin is not None:
            inside = [(ts, ln) for ts, ln in lines if t_begin <= ts <= t_end]
            lines = inside or [(ts, ln) for ts, ln in lines if t_begin - 0.1 <= ts <= t_end + 0.1]
        for ts, ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}

    def stop(self, t_begin=None, t_end=None):
        out = self.window(t_begin, t_end)
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        return out


def workload_string(name):
    """The SAME string in both arms (ours and --impl reference), so the driver's same_config check can hold."""
    api, b, h, n, d, causal, dtype_name = WORKLOADS[name]
    return f"{name}: {api} B={b} H={h} N={n} D={d} causal={causal} {dtype_name}"


def measure_ceilings(dev):
    """Dense INT8 and FP8 GEMM ceilings of THIS GPU, measured in this run (torch._int_mm and
    torch._scaled_mm at 8192^3, best of 10, CUDA events), next to a bf16 matmul taken the same way.  These are
    library GEMMs used as roofline denominators only; nothing on the product path calls them."""
    import torch
    n = 8192
    out = {"shape": f"{n}^3", "how": "best of 10 after 3 warm-ups, CUDA events, 2*N^3 ops"}

    def best(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b_.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b_))
        return 2.0 * n ** 3 / (min(ts) * 1e-3) * 1e-12

    try:
        x = torch.randn(n, n, device=dev, dtype=torch.bfloat16)
        y = torch.randn(n, n, device=dev, dtype=torch.bfloat16)
        out["bf16_tflops"] = best(lambda: torch.matmul(x, y))
    except Exception as e:  # pragma: no cover
        out["bf16_tflops"], out["bf16_error"] = None, f"{type(e).__name__}: {e}"[:160]
    try:
        a8 = torch.randint(-100, 100, (n, n), device=dev, dtype=torch.int8)
        b8 = torch.randint(-100, 100, (n, n), device=dev, dtype=torch.int8).t()   # column-major B, as cuBLASLt wants
        out["int8_tops"] = best(lambda: torch._int_mm(a8, b8))
    except Exception as e:
        out["int8_tops"], out["int8_error"] = None, f"{type(e).__name__}: {e}"[:160]
    try:
        af = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn)
        bf = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn).t()
        one = torch.ones((), device=dev, dtype=torch.float32)
        out["fp8_tflops"] = best(lambda: torch._scaled_mm(af, bf, scale_a=one, scale_b=one, out_dtype=torch.bfloat16))
    except Exception as e:
        out["fp8_tflops"], out["fp8_error"] = None, f"{type(e).__name__}: {e}"[:160]
    return out


def mixed_ceiling(qk_peak, pv_peak):
    """Half of the attention flops run as QK^T (one tensor rate), half as P.V (another): the time-weighted
    ceiling is the harmonic combination."""
    if not qk_peak or not pv_peak:
        return None
    return 2.0 / (1.0 / qk_peak + 1.0 / pv_peak)


def run_sweep(dev, sampler, ceilings, reps=5):
    """The metric's whole range in one run: hd {64,128} x {full, causal} x N {1K..32K} x PV {fp8, f16}, B=4 H=32
    (protocol of /root/reference/bench/bench_qk_int8_pv_fp8_cuda.py:70-98: random INT8 Q/K codes, random scales;
    kernel-only = the attention launch alone, median of `reps` CUDA-event timings of single launches through the
    C ABI).  `op_tops` = the WHOLE OPERATOR through the public Python API on fp16/bf16 q, k, v (quantisation
    pre-pass + kernel + host cost of the call: back-to-back API calls, no CUDA-graph replay).  Every entry carries
    the SM clock samples taken while it ran."""
    import torch
    import sageattention_b200 as sb
    from sageattention_b200 import core
    B, H = 4, 32
    out = []
    g = torch.Generator(device=dev).manual_seed(0)

    def timed(fn, n_rep):
        evs = []
        for _ in range(n_rep):
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b_.record()
            evs.append((a, b_))
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b_) for a, b_ in evs)
        return ts[len(ts) // 2]

    for d in (128, 64):
        for n in (1024, 2048, 4096, 8192, 16384, 32768):
            q8 = torch.randint(-95, 95, (B, H, n, d), dtype=torch.int8, device=dev, generator=g)
            k8 = torch.randint(-95, 95, (B, H, n, d), dtype=torch.int8, device=dev, generator=g)
            qs = torch.rand(B, H, (n + 127) // 128 * 32, device=dev, generator=g) * 0.01 + 0.005
            ks = torch.rand(B, H, (n + 63) // 64 * 4, device=dev, generator=g) * 0.01 + 0.005
            v16 = torch.randn(B, H, n, d, device=dev, generator=g).to(torch.float16)
            v8 = torch.randn(B, H, d, (n + 127) // 128 * 128, device=dev, generator=g).to(torch.float8_e4m3fn)
            vs = torch.rand(B, H, d, device=dev, generator=g) + 0.5
            o = torch.empty(B, H, n, d, dtype=torch.float16, device=dev)
            qf = torch.randn(B, H, n, d, device=dev, generator=g).to(torch.float16)
            kf = torch.randn(B, H, n, d, device=dev, generator=g).to(torch.float16)
            sm = d ** -0.5
            for causal in (False, True):
                flops = flops_of(B, H, n, d, causal)
                for pv in ("fp8", "f16"):
                    fp8 = pv == "fp8"
                    kern = lambda: core._launch(fp8, q8, k8, v8 if fp8 else v16, o, qs, ks, vs if fp8 else None, None,
                                                "HND", causal, 3, 32, sm, False)
                    api = sb.sageattn_qk_int8_pv_fp8_cuda if fp8 else sb.sageattn_qk_int8_pv_fp16_cuda
                    op = lambda: api(qf, kf, v16, is_causal=causal)
                    n_rep = reps if n >= 8192 else 4 * reps
                    for _ in range(3):
                        kern()
                        op()
                    torch.cuda.synchronize()
                    t0 = time.monotonic()
                    k_ms = timed(kern, n_rep)
                    # the operator: n_rep back-to-back API calls between two events (what a caller's loop sees)
                    a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    for _ in range(n_rep):
                        op()
                    b_.record()
                    torch.cuda.synchronize()
                    op_ms = a.elapsed_time(b_) / n_rep
                    t1 = time.monotonic()
                    ck = sampler.window(t0, t1) if sampler is not None else None
                    pv_peak = ceilings.get("fp8_tflops") if fp8 else ceilings.get("bf16_tflops")
                    ceil = mixed_ceiling(ceilings.get("int8_tops"), pv_peak)
                    k_tops = flops / (k_ms * 1e-3) * 1e-12
                    out.append({"hd": d, "n": n, "causal": causal, "pv": pv, "kernel_ms": round(k_ms, 4),
                                "kernel_tops": round(k_tops, 1), "op_ms": round(op_ms, 4),
                                "op_tops": round(flops / (op_ms * 1e-3) * 1e-12, 1),
                                "frac_of_measured_ceiling": round(k_tops / ceil, 4) if ceil else None,
                                "clocks": ck})
            del q8, k8, qs, ks, v16, v8, vs, o, qf, kf
            torch.cuda.empty_cache()
    return out


def run_sp(dev, world, rank, barrier, max_over_ranks, transport="nccl", chunks=1, comm_sms=0, reps=5):
    """The sequence-parallel DiT record (BASELINE.json config 5): one Wan2.1-14B attention layer, B1 H40 N75600 hd128
    bf16, through `sageattn`.  World 1: the layer on one GPU.  World P: activations sharded along the sequence
    ([1, N/P, H, D] per rank), `ulysses_sageattn` = all-to-all in (scatter heads, gather sequence), attention on H/P
    heads of the full sequence, all-to-all out; `comm_ms` = the same call with a null attention function, i.e. the
    two all-to-alls and the gather / interleave copies alone; `one_gpu_ms` = the whole layer timed on rank 0's GPU in
    this same run, `efficiency` = one_gpu_ms / (P * ms).  Device time, max over ranks, barrier on both sides."""
    import torch
    import sageattention_b200 as sb
    from sageattention_b200 import parallel
    b, h, n, d = 1, 40, 75600, 128
    flops = flops_of(b, h, n, d, False)
    g = torch.Generator(device=dev).manual_seed(100 + rank)

    def timed(fn):
        for _ in range(3):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1) / reps)

    q, k, v = (torch.randn(b, h, n, d, device=dev, generator=g).to(torch.bfloat16) for _ in range(3))
    one_ms = timed(lambda: sb.sageattn(q, k, v, tensor_layout="HND", is_causal=False))
    rec = {"layer": f"Wan2.1-14B attention layer: sageattn B={b} H={h} N={n} D={d} bfloat16 non-causal",
           "one_gpu_ms": round(one_ms, 3), "one_gpu_tops": round(flops / (one_ms * 1e-3) * 1e-12, 1)}
    del q, k, v
    if world == 1:
        rec.update({"mode": "single GPU", "ms": rec["one_gpu_ms"], "tops": rec["one_gpu_tops"], "comm_ms": 0.0,
                    "efficiency": 1.0})
        return rec
    assert h % world == 0 and n % world == 0
    n_loc = n // world
    qs, ks_, vs_ = (torch.randn(b, n_loc, h, d, device=dev, generator=g).to(torch.bfloat16) for _ in range(3))
    full = lambda: parallel.ulysses_sageattn(qs, ks_, vs_, attn_fn=sb.sageattn, transport=transport, head_chunks=chunks,
                                             comm_sms=comm_sms, is_causal=False)
    null_attn = lambda q_, k_, v_, **kw: torch.empty_like(q_)
    comm = lambda: parallel.ulysses_sageattn(qs, ks_, vs_, attn_fn=null_attn, transport=transport, is_causal=False)
    ms = timed(full)
    comm_ms = timed(comm)
    rec.update({"mode": f"Ulysses all-to-all ({transport}) over {world} GPUs, sequence shard {n_loc} tokens per rank, "
                        f"{chunks} head group(s) pipelined, {comm_sms} SMs left to the collective",
                "ms": round(ms, 3), "tops": round(flops / (ms * 1e-3) * 1e-12, 1), "comm_ms": round(comm_ms, 3),
                "efficiency": round(one_ms / (world * ms), 4),
                "a2a_bytes_per_rank": int(4 * b * n_loc * h * d * 2 * (world - 1) / world)})
    # the same layer with the quantised payload: pre-pass on the sequence shards, INT8 / e4m3 codes and
    # scales through the inbound all-to-alls (1 byte per element instead of 2), 16-bit output back
    try:
        class _NullAttention(parallel._Q8CudaOps):
            def attention(self, q8, k8, v8, q_scale, k_scale, v_scale, out_dtype, is_causal, sm_scale, o_peers=None):
                return torch.empty((1, q8.size(1), q8.size(2), q8.size(3)), dtype=out_dtype, device=q8.device)
        full8 = lambda: parallel.ulysses_sageattn(qs, ks_, vs_, payload="int8", is_causal=False)
        comm8 = lambda: parallel.ulysses_sageattn(qs, ks_, vs_, payload="int8", _q8_ops=_NullAttention(), is_causal=False)
        ms8 = timed(full8)
        comm8_ms = timed(comm8)
        m_rows = (n_loc + 127) // 128 * 128
        rec["int8_payload"] = {
            "ms": round(ms8, 3), "tops": round(flops / (ms8 * 1e-3) * 1e-12, 1),
            "comm_and_prepass_ms": round(comm8_ms, 3), "efficiency": round(one_ms / (world * ms8), 4),
            "a2a_bytes_per_rank": int((3 * m_rows + 2 * n_loc) * b * h * d * (world - 1) / world),
            "note": "comm_and_prepass_ms = the call with a null attention kernel: statistics, halo exchange, quantisers, "
                    "both inbound all-to-alls, re-assembly, outbound all-to-all (the 16-bit path's comm_ms excludes its "
                    "pre-pass, which runs inside the attention call there)"}
    except Exception as e:  # the record of the default path must survive a failure here
        rec["int8_payload"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    # and with the quantisers storing straight into the peers' symmetric-memory arenas over NVLink: no inbound collective
    try:
        full9 = lambda: parallel.ulysses_sageattn(qs, ks_, vs_, payload="int8", transport="p2p", is_causal=False)
        comm9 = lambda: parallel.ulysses_sageattn(qs, ks_, vs_, payload="int8", transport="p2p", _q8_ops=_NullAttention(),
                                                  is_causal=False)
        ms9 = timed(full9)
        comm9_ms = timed(comm9)
        rec["int8_payload_peer_stores"] = {
            "ms": round(ms9, 3), "tops": round(flops / (ms9 * 1e-3) * 1e-12, 1),
            "comm_and_prepass_ms": round(comm9_ms, 3), "efficiency": round(one_ms / (world * ms9), 4),
            "note": "quantiser launches write INT8 / e4m3 codes and scales into the destination ranks' arenas (one base "
                    "pointer per destination) and the attention kernel's TMA-store epilogue writes every destination's "
                    "output rows into its arena; three signal-pad barriers; only the statistics all-gather and the "
                    "halo go through NCCL"}
    except Exception as e:
        rec["int8_payload_peer_stores"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    cands = [("16-bit payload, NCCL all-to-alls (ulysses_sageattn defaults)", rec["ms"], rec["efficiency"])]
    for key, label in (("int8_payload", 'payload="int8"'), ("int8_payload_peer_stores", 'payload="int8", transport="p2p"')):
        if "ms" in rec.get(key, {}):
            cands.append((label, rec[key]["ms"], rec[key]["efficiency"]))
    best = min(cands, key=lambda c: c[1])
    rec["best"] = {"variant": best[0], "ms": best[1], "efficiency": best[2]}
    return rec


def cpu_sample(api, b, h, n, d, causal, dtype_name, budget_s=20.0):
    """The CPU restatement of the reference algorithm (oracle/sage_oracle.py) on a bounded sample:
    one (batch, head-group) slice of the workload, all host threads."""
    import torch
    from oracle import sage_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    dtype = getattr(torch, dtype_name)
    pv = "f16" if "fp16" in api else "f8"
    hs = min(h, 8)
    ns = min(n, 4096)
    g = torch.Generator().manual_seed(0)
    q = torch.randn(1, hs, ns, d, generator=g).to(dtype)
    k = torch.randn(1, hs, ns, d, generator=g).to(dtype)
    v = torch.randn(1, hs, ns, d, generator=g).to(dtype)
    kw = {"pv_accum_dtype": "fp32"} if pv == "f16" else {"pv_accum_dtype": "fp32+fp32"}
    t0 = time.perf_counter()
    reps = 0
    while True:
        O.sageattn_ref(q, k, v, pv=pv, is_causal=causal, **kw)
        reps += 1
        el = time.perf_counter() - t0
        if el > budget_s or reps >= 20:
            break
    t = el / reps
    tops = flops_of(1, hs, ns, d, causal) / t * 1e-12
    sample = f"B=1 H={hs} N={ns} D={d} slice of the workload, {reps} rep(s), {t:.2f} s each"
    return tops, sample, t


def run_reference(args):
    api, b, h, n, d, causal, dtype_name = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, args.steps)
    tops_l, sample, ts = [], "", []
    # every step is a bounded sample (about 3.5 s of CPU work: the driver's --steps 20 --warmup 5 takes 90 s); the arm
    # stops after about two and a half minutes of wall time whatever --steps says (the GPU arm's default is 400 steps)
    # and reports how many steps it timed
    t_wall = time.time()
    warm = args.warmup
    for i in range(warm + steps):
        tops, sample, t = cpu_sample(api, b, h, n, d, causal, dtype_name, budget_s=4.0)
        if i >= warm:
            tops_l.append(tops)
            ts.append(t)
        if time.time() - t_wall > 150 and tops_l:
            break
    val = statistics.mean(tops_l)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "TOPS", "n_gpus": args.gpus,
        "steps": len(tops_l), "warmup": warm, "ms_per_step": statistics.mean(ts) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8+fp16" if "fp16" in api else "int8+fp8",
        "data": "synthetic",
        "config": {"workload": workload_string(args.workload)},
        "cpu_baseline": {"value": val, "unit": "TOPS", "cores": os.cpu_count(), "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "TOPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference has no CPU implementation of this path; this is the CPU restatement (oracle/) of its algorithm",
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2_fp16", choices=sorted(WORKLOADS))
    ap.add_argument("--no-sweep", action="store_true", help="skip the hd x mask x N x PV sweep (N=1 only) in the line")
    ap.add_argument("--no-sp", action="store_true", help="skip the sequence-parallel Wan layer record")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--sp", default="heads", choices=["heads", "ulysses"],
                    help="DiT workloads under torchrun: head sharding (no collective) or Ulysses all-to-all")
    ap.add_argument("--sp-chunks", type=int, default=1,
                    help="Ulysses: groups of heads whose all-to-alls overlap the attention of another group")
    ap.add_argument("--sp-comm-sms", type=int, default=0,
                    help="Ulysses with --sp-chunks > 1: SMs the attention grid leaves free for the NCCL kernels")
    ap.add_argument("--sp-transport", default="nccl", choices=["nccl", "p2p"],
                    help="Ulysses: NCCL all-to-all, or copy-engine pushes into the peers' symmetric memory")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist
    import sageattention_b200 as sb
    from sageattention_b200 import _lib, core

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W = max(3, args.warmup)
    K = max(1, args.steps)

    api, b, h, n, d, causal, dtype_name = WORKLOADS[args.workload]
    dtype = getattr(torch, dtype_name)
    fn = getattr(sb, api)
    sharded = args.workload in ("hunyuan", "wan") and world > 1
    if sharded:
        assert h % world == 0, "heads must divide across ranks"
        h_loc = h // world
        scaling = "strong"
    else:
        h_loc = h
        scaling = "weak"
    torch.manual_seed(rank)
    ulysses = sharded and args.sp == "ulysses"
    if ulysses:
        # sequence-sharded activations [B, N/P, H, D]; the sequence is padded to a multiple of P
        from sageattention_b200 import parallel
        n_loc = (n + world - 1) // world
        shape = (b, n_loc, h, d)
        base_fn = fn
        fn = lambda q_, k_, v_, is_causal=False: parallel.ulysses_sageattn(q_, k_, v_, attn_fn=base_fn,
                                                                           head_chunks=args.sp_chunks,
                                                                           transport=args.sp_transport,
                                                                           is_causal=is_causal)
        n = n_loc * world
    else:
        shape = (b, h_loc, n, d)
    q = torch.randn(shape, dtype=dtype, device=dev)
    k = torch.randn(shape, dtype=dtype, device=dev)
    v = torch.randn(shape, dtype=dtype, device=dev)
    flops_rank = flops_of(b, h_loc, n, d, causal)
    flops_job = flops_rank * world

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident timing (value) + kernel-only events (roofline)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # started before the warm-up so that it is streaming when the timed region begins
    for _ in range(W):
        fn(q, k, v, is_causal=causal)
    barrier()
    if rank == 0:
        sampler.wait_first()
    core.KERNEL_EVENTS = []
    launches0 = _lib.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_begin = time.monotonic()
    e0.record()
    for _ in range(K):
        o = fn(q, k, v, is_causal=causal)
    e1.record()
    barrier()
    t_end = time.monotonic()
    launches = _lib.LAUNCHES - launches0
    t_ms = max_over_ranks(e0.elapsed_time(e1)) / K
    kern_ms = [a.elapsed_time(bb) for a, bb in core.KERNEL_EVENTS]
    core.KERNEL_EVENTS = None
    kern_ms_avg = max_over_ranks(sum(kern_ms) / len(kern_ms))
    clocks = sampler.window(t_begin, t_end) if rank == 0 else None
    value = flops_job / (t_ms * 1e-3) * 1e-12

    # ---------------- end to end with host buffers
    qh, kh, vh = (torch.randn(shape, dtype=dtype).pin_memory() for _ in range(3))
    oh = torch.empty(shape, dtype=dtype).pin_memory()
    ohs = [oh, torch.empty(shape, dtype=dtype).pin_memory()]
    from sageattention_b200.host import HostPipeline
    pipe = HostPipeline(fn, shape, dtype, device=dev)

    def e2e_step(i):
        # every step: H2D of q, k, v from pinned host memory, the operator, D2H of the output;
        # the pipeline overlaps step i's copies with its neighbours' kernels (three streams)
        pipe.submit(qh, kh, vh, ohs[i % 2], is_causal=causal)

    ke = max(4, min(K, 30))
    for i in range(6):
        e2e_step(i)
    pipe.drain()
    barrier()
    e0.record()
    pipe.fork()
    for i in range(ke):
        e2e_step(i)
    pipe.drain()
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1)) / ke
    e2e_val = flops_job / (e2e_ms * 1e-3) * 1e-12
    nbytes = q.numel() * q.element_size()
    del pipe, qh, kh, vh, oh, ohs
    sp = None
    if not args.no_sp and not sharded:
        del q, k, v, o
        torch.cuda.empty_cache()
        try:
            sp = run_sp(dev, world, rank, barrier, max_over_ranks, transport=args.sp_transport, chunks=args.sp_chunks,
                        comm_sms=args.sp_comm_sms)
        except Exception as e:  # the headline line must survive a failure here
            sp = {"error": f"{type(e).__name__}: {e}"[:300]}

    if rank != 0:
        sampler.stop()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    bf16_peak, hbm_peak, basis = peaks()
    ceilings = measure_ceilings(dev)
    fp8 = "fp16_cuda" not in api
    ceil_meas = mixed_ceiling(ceilings.get("int8_tops"), ceilings.get("fp8_tflops") if fp8 else ceilings.get("bf16_tflops"))
    mix = 2.0 if fp8 else 4.0 / 3.0
    peak = bf16_peak * mix
    achieved = flops_rank / (kern_ms_avg * 1e-3) * 1e-12
    line = {
        "metric": METRIC, "value": value, "unit": "TOPS", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": t_ms, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "int8+fp8" if fp8 else "int8+fp16", "data": "synthetic",
        "config": {"workload": workload_string(args.workload), "heads_per_rank": h_loc,
                   "split": (f"Ulysses all-to-all ({args.sp_transport}), {args.sp_chunks} head group(s) pipelined" if ulysses
                             else ("heads across ranks, no collective" if sharded else "independent replicas per rank, no collective")),
                   "l2": "inputs+outputs per step (%.0f MB) exceed the 126 MB L2" % (4 * nbytes / 1e6)},
        "e2e": {"value": e2e_val, "unit": "TOPS", "h2d_bytes_per_step": 3 * nbytes * world,
                "d2h_bytes_per_step": nbytes * world, "ms_per_step": e2e_ms},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": NCU_TRAFFIC.get(args.workload),
                     "traffic_source": NCU_TRAFFIC_SOURCE.get(args.workload), "kernel": "sage_attn_fwd_kernel",
                     "kernel_ms": kern_ms_avg,
                     "peak_basis": f"{basis} bf16 cuBLAS {bf16_peak:.0f} TF/s x {mix:.3f} (i8/fp8 MMAs run at 2x the f16 rate); "
                                   f"nominal {'4500' if fp8 else '3000'}",
                     "frac_of_nominal": achieved / (4500.0 if fp8 else 3000.0),
                     "measured_ceilings": ceilings, "peak_from_measured_int8_and_pv_gemm": ceil_meas,
                     "frac_of_measured_int8_and_pv_gemm": (achieved / ceil_meas) if ceil_meas else None},
        "kernel_only_tops": achieved,
    }
    if sp is not None:
        line["sp"] = sp
    if world == 1 and not args.no_sweep:
        try:
            line["sweep"] = run_sweep(dev, sampler, ceilings)
        except Exception as e:
            line["sweep"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    sampler.stop()
    if not args.no_cpu and world == 1:
        ctops, sample, _ = cpu_sample(api, b, h, n, d, causal, dtype_name)
        line["cpu_baseline"] = {"value": ctops, "unit": "TOPS", "cores": os.cpu_count(), "kind": "port",
                                "sample": sample}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
