"""Run tools/bench_kernel.py and a correctness check for several library variants (tools/_bin/libsage_*.so),
each in its own process (development tool, run on the GPU box):
    python tools/sweep_variants.py base,v1,v2 --pv f8 --d 128 --n 8192 --causal 0
Output: gpurun_out/variants.jsonl (one line per variant x configuration, with the variant name) and a
max-abs / cosine check of every variant's output against the FIRST variant on fixed inputs."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
names = sys.argv[1].split(",")
rest = sys.argv[2:]
out = os.path.join(ROOT, "gpurun_out", "variants.jsonl")
os.makedirs(os.path.dirname(out), exist_ok=True)

CHECK = r'''
import sys, torch
sys.path.insert(0, %r)
from sageattention_b200 import core
torch.manual_seed(1)
dev = "cuda:0"
res = {}
for pv in ("f8", "f16"):
    for d in (128, 64):
        for causal in (False, True):
            n = 1000
            q8 = torch.randint(-95, 95, (2, 4, n, d), dtype=torch.int8, device=dev)
            k8 = torch.randint(-95, 95, (2, 4, n, d), dtype=torch.int8, device=dev)
            qs = torch.rand(2, 4, (n + 127) // 128 * 32, device=dev) * 0.01 + 0.005
            ks = torch.rand(2, 4, (n + 63) // 64 * 4, device=dev) * 0.01 + 0.005
            if pv == "f8":
                v = torch.randn(2, 4, d, 1024, device=dev).to(torch.float8_e4m3fn); vs = torch.rand(2, 4, d, device=dev) + 0.5
            else:
                v = torch.randn(2, 4, n, d, device=dev).half(); vs = None
            o = torch.empty(2, 4, n, d, dtype=torch.float16, device=dev)
            lse = core._launch(pv == "f8", q8, k8, v, o, qs, ks, vs, None, "HND", causal, 3, 32, d ** -0.5, True)
            torch.cuda.synchronize()
            res[(pv, d, causal)] = (o.cpu(), lse.cpu())
torch.save(res, sys.argv[1])
'''

ref = None
for name in names:
    lib = os.path.join(ROOT, "tools", "_bin", f"libsage_{name}.so") if name != "prod" else \
        os.path.join(ROOT, "sageattention_b200", "libsage_b200.so")
    env = dict(os.environ, SAGE_B200_LIB=lib)
    pt = os.path.join(ROOT, "gpurun_out", f"check_{name}.pt")
    r = subprocess.run([sys.executable, "-c", CHECK % ROOT, pt], env=env, capture_output=True, text=True, timeout=600)
    if r.returncode != 0:
        print(f"[{name}] CHECK FAILED TO RUN\n{r.stderr[-2000:]}", flush=True)
        continue
    import torch
    cur = torch.load(pt)
    os.remove(pt)
    if ref is None:
        ref = cur
        print(f"[{name}] reference variant for the output check", flush=True)
    else:
        worst = 1.0
        dl = 0.0
        do = 0.0
        for key, (o, lse) in cur.items():
            ro, rl = ref[key]
            do = max(do, (o.float() - ro.float()).abs().max().item())
            cos = torch.nn.functional.cosine_similarity(o.float().flatten(), ro.float().flatten(), dim=0).item()
            worst = min(worst, cos)
            dl = max(dl, (lse - rl).abs().max().item())
            if not torch.isfinite(o.float()).all():
                worst = float("nan")
        print(f"[{name}] vs {names[0]}: worst cos {worst:.7f}  max |dO| {do:.2e}  max |dLSE| {dl:.2e}", flush=True)
    tmp = os.path.join(ROOT, "gpurun_out", f"_sweep_{name}.jsonl")
    if os.path.exists(tmp):
        os.remove(tmp)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "bench_kernel.py"), "--out", tmp] + rest, env=env,
                       capture_output=True, text=True, timeout=1800)
    if r.returncode != 0:
        print(f"[{name}] BENCH FAILED\n{r.stderr[-2000:]}", flush=True)
        continue
    with open(out, "a") as f:
        for line in open(tmp):
            rec = json.loads(line)
            rec["variant"] = name
            f.write(json.dumps(rec) + "\n")
            print(f"[{name}] {rec['pv']} d{rec['d']} causal={int(rec['causal'])} n={rec['n']}: {rec['tops']:.0f} TOPS "
                  f"(best {rec['tops_best']:.0f})", flush=True)
    os.remove(tmp)
