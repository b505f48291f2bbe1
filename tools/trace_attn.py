"""Pipeline timeline of CTA 0 of the attention kernel across its first work items (development tool).
    python tools/build_flags.py trace --only attn_f16_sm100.cu -DSAGE_TRACE     (attn_f8_sm100.cu for the e4m3 flavour)
    SAGE_B200_LIB=tools/_bin/libsage_trace.so python tools/trace_attn.py [f8|f16] [head_dim] [N] [causal]
Prints, per KV step (counted across items), SM-clock intervals of the two softmax warpgroups."""
import ctypes, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sageattention_b200 as S
from sageattention_b200 import _lib

kind = sys.argv[1] if len(sys.argv) > 1 else "f8"
D = int(sys.argv[2]) if len(sys.argv) > 2 else 128
N = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
causal = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
torch.manual_seed(0)
B, H = 4, 32
q, k, v = (torch.randn(B, H, N, D, device="cuda", dtype=torch.float16) for _ in range(3))
fn = S.sageattn_qk_int8_pv_fp8_cuda if kind == "f8" else S.sageattn_qk_int8_pv_fp16_cuda
for _ in range(3):
    fn(q, k, v, tensor_layout="HND", is_causal=causal)
torch.cuda.synchronize()
lib = ctypes.CDLL(_lib._LIB_PATH)
steps, slots = 64, 12
buf = (ctypes.c_longlong * (3 * steps * slots))()
rc = lib.sage_debug_trace(buf)
assert rc == 0, rc
t = np.frombuffer(buf, dtype=np.int64).reshape(3, steps, slots).copy()
t0 = t[0, 0, 0]
t -= t0
print(f"{kind} D={D} N={N} causal={causal}")
spec = bool(int(os.environ.get("TRACE_SPEC", "0")))   # stamps of softmax_steps_spec (SAGE_SPEC)
for j in range(0, 40):
    for g in range(2):
        a = t[g, j]
        if spec:
            print(f"j={j:2d} WG{g}: start {a[0]:7d} upper half (ld, max, exp) {a[6]-a[0]:5d} idle for lower S {a[1]-a[6]:5d} ld {a[2]-a[1]:4d} "
                  f"max + decision + publish upper {a[5]-a[2]:5d} exp lower {a[7]-a[5]:5d} publish lower {a[4]-a[7]:4d} | period {t[g, j+1, 0]-a[0]:5d}")
            continue
        ep = f" | setup {a[10]-a[4]:5d} PVwait {a[8]-a[10]:5d} bar {a[11]-a[8]:5d} epilogue {a[9]-a[11]:5d}" if a[9] > 0 else ""
        m = t[2, j]
        ep += f" | MMA: wait_begin {m[g*4]-a[4]:6d} P_seen {m[g*4+1]-a[4]:5d} pv_issued {m[g*4+2]-a[4]:5d} (rel. to P arrive)"
        print(f"j={j:2d} WG{g}: start {a[0]:7d} wait {a[1]-a[0]:5d} ld {a[2]-a[1]:4d} max {a[5]-a[2]:4d} turn {a[6]-a[5]:5d} exp {a[7]-a[6]:5d} pack {a[3]-a[7]:4d} store+arrive {a[4]-a[3]:4d} | period {t[g, j+1, 0]-a[0]:5d}{ep}")
per = (t[0, 56, 0] - t[0, 8, 0]) / 48
print(f"mean period {per:.0f} cycles per step over steps 8..56")
