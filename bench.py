"""bench.py - the measurement contract for the SageAttention hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME] [--no-sweep]
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
    "c3_fp8": ("sageattn_qk_int8_pv_fp8_cuda", 4, 32, 4096, 128, False, "float16"),
    "c3_fp8_8k": ("sageattn_qk_int8_pv_fp8_cuda", 4, 32, 8192, 128, False, "bfloat16"),
    "hunyuan": ("sageattn", 1, 24, 119056, 128, False, "bfloat16"),
    "wan": ("sageattn", 1, 40, 75600, 128, False, "bfloat16"),
}
# dram__bytes_read.sum + dram__bytes_write.sum of the attention kernel, one `ncu --set full` capture
# per workload (profiles/r1b_ncu_attn_f16_c2_summary.txt, r1b_ncu_attn_f8_c3_summary.txt: the persistent
# kernels); algorithmic bytes at C2 fp16 = 402.7e6
NCU_TRAFFIC = {"c2_fp16": 377.14e6, "c3_fp8": 302.79e6}
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

    def stop(self, t_begin=None, t_end=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # only samples taken inside the timed region count; a region shorter than the 50 ms sampling
        # period falls back to the samples within 100 ms of it
        lines = self.lines
        if t_begin is not None:
            lines = [(ts, ln) for ts, ln in self.lines if t_begin <= ts <= t_end]
            if not lines:
                lines = [(ts, ln) for ts, ln in self.lines if t_begin - 0.1 <= ts <= t_end + 0.1]
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
    for i in range(args.warmup + steps):
        tops, sample, t = cpu_sample(api, b, h, n, d, causal, dtype_name, budget_s=8.0)
        if i >= args.warmup:
            tops_l.append(tops)
            ts.append(t)
        if sum(ts) > 150:
            break
    val = statistics.mean(tops_l)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "TOPS", "n_gpus": args.gpus,
        "steps": len(tops_l), "warmup": args.warmup, "ms_per_step": statistics.mean(ts) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8+fp16" if "fp16" in api else "int8+fp8",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: {api} B={b} H={h} N={n} D={d} causal={causal} {dtype_name}"},
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
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--sp", default="heads", choices=["heads", "ulysses"],
                    help="DiT workloads under torchrun: head sharding (no collective) or Ulysses all-to-all")
    ap.add_argument("--sp-chunks", type=int, default=1,
                    help="Ulysses: groups of heads whose all-to-alls overlap the attention of another group")
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
    clocks = sampler.stop(t_begin, t_end) if rank == 0 else None
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

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    bf16_peak, hbm_peak, basis = peaks()
    fp8 = "fp16_cuda" not in api
    mix = 2.0 if fp8 else 4.0 / 3.0
    peak = bf16_peak * mix
    achieved = flops_rank / (kern_ms_avg * 1e-3) * 1e-12
    line = {
        "metric": METRIC, "value": value, "unit": "TOPS", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": t_ms, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "int8+fp8" if fp8 else "int8+fp16", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {api} B={b} H={h} N={n} D={d} causal={causal} {dtype_name}, "
                               f"heads per rank {h_loc}" + (f", Ulysses all-to-all ({args.sp_transport}), {args.sp_chunks} head group(s) pipelined" if ulysses else ""),
                   "l2": "inputs+outputs per step (%.0f MB) exceed the 126 MB L2" % (4 * nbytes / 1e6)},
        "e2e": {"value": e2e_val, "unit": "TOPS", "h2d_bytes_per_step": 3 * nbytes * world,
                "d2h_bytes_per_step": nbytes * world, "ms_per_step": e2e_ms},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": NCU_TRAFFIC.get(args.workload), "kernel": "sage_attn_fwd_kernel",
                     "kernel_ms": kern_ms_avg,
                     "peak_basis": f"{basis} bf16 cuBLAS {bf16_peak:.0f} TF/s x {mix:.3f} (i8/fp8 MMAs run at 2x the f16 rate); "
                                   f"nominal {'4500' if fp8 else '3000'}",
                     "frac_of_nominal": achieved / (4500.0 if fp8 else 3000.0)},
        "kernel_only_tops": achieved,
    }
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
