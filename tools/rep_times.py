"""Per-repetition device times of one attention kernel configuration, in launch order (development tool): tells a
clock / power drift (monotone) from a kernel that settles into different regimes from launch to launch (scatter).
    SAGE_B200_LIB=tools/_bin/libsage_X.so [REP_VERBOSE=1] python tools/rep_times.py [f8|f16] [head_dim] [N] [causal] [reps]
On this pool a B200 runs the first ~25 ms of a kernel train at full clocks and then settles at its power limit, so a faster
kernel shows a larger gain in `burst` than in `sustained`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sageattention_b200 import _lib, core  # noqa: E402

pv = sys.argv[1] if len(sys.argv) > 1 else "f16"
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
n = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
causal = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 60
b, h = 4, 32
dev = "cuda:0"
torch.manual_seed(0)
q8 = torch.randint(-95, 95, (b, h, n, d), dtype=torch.int8, device=dev)
k8 = torch.randint(-95, 95, (b, h, n, d), dtype=torch.int8, device=dev)
qs = torch.rand(b, h, (n + 127) // 128 * 32, device=dev) * 0.01 + 0.005
ks = torch.rand(b, h, (n + 63) // 64 * 4, device=dev) * 0.01 + 0.005
if pv == "f8":
    v = torch.randn(b, h, d, (n + 127) // 128 * 128, device=dev).to(torch.float8_e4m3fn)
    vs = torch.rand(b, h, d, device=dev) + 0.5
else:
    v = torch.randn(b, h, n, d, device=dev).half()
    vs = None
o = torch.empty(b, h, n, d, dtype=torch.float16, device=dev)


def fn():
    core._launch(pv == "f8", q8, k8, v, o, qs, ks, vs, None, "HND", causal, _lib.GRAN_PER_THREAD, 32, d ** -0.5, False)


for _ in range(10):
    fn()
torch.cuda.synchronize()
evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
for a, e in evs:
    a.record()
    fn()
    e.record()
torch.cuda.synchronize()
ts = [a.elapsed_time(e) for a, e in evs]
flops = 4.0 * b * h * n * n * d * (0.5 if causal else 1.0)
name = os.path.basename(os.environ.get("SAGE_B200_LIB", "prod"))
med = lambda x: sorted(x)[len(x) // 2]
if os.environ.get("REP_VERBOSE"):
    print(" ".join(f"{flops / t * 1e-9:.0f}" for t in ts))
# burst: the launches before the power limit bites (about 25 ms of work); sustained: the second half of the run
nb = max(2, min(len(ts) // 4, int(25.0 / med(ts))))
print(f"{name:22s} {pv} d{d} n={n:6d} causal={int(causal)}: burst {flops / med(ts[1:1 + nb]) * 1e-9:5.0f}  "
      f"sustained {flops / med(ts[len(ts) // 2:]) * 1e-9:5.0f}  best {flops / min(ts) * 1e-9:5.0f} TOPS ({len(ts)} launches in a row)")
