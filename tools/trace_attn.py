"""Pipeline timeline of one CTA of the attention kernel (development tool).
    python tools/build_variant.py trace sageattention_b200/csrc/attn_sm100.cu -DSAGE_TRACE
    SAGE_B200_LIB=tools/_bin/libsage_trace.so python tools/trace_attn.py [f8|f16] [head_dim]
Prints, per KV step, SM-clock intervals of the two softmax warpgroups and of the MMA issuer."""
import ctypes, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sageattention_b200 as S
from sageattention_b200 import _lib

kind = sys.argv[1] if len(sys.argv) > 1 else "f8"
D = int(sys.argv[2]) if len(sys.argv) > 2 else 128
torch.manual_seed(0)
B, H, N = 2, 8, 8192
q, k, v = (torch.randn(B, H, N, D, device="cuda", dtype=torch.float16) for _ in range(3))
fn = S.sageattn_qk_int8_pv_fp8_cuda if kind == "f8" else S.sageattn_qk_int8_pv_fp16_cuda
for _ in range(3):
    fn(q, k, v, tensor_layout="HND", is_causal=False)
torch.cuda.synchronize()
lib = ctypes.CDLL(_lib._LIB_PATH)
steps, slots = 48, 8
buf = (ctypes.c_longlong * (3 * steps * slots))()
rc = lib.sage_debug_trace(buf)
assert rc == 0, rc
t = np.frombuffer(buf, dtype=np.int64).reshape(3, steps, slots).copy()
t0 = t[0, 0, 0]
t -= t0
print(f"{kind} D={D}: per step  WG: wait_begin | S ready (+wait) | ld done | softmax done | P arrived ;  MMA t: wait_begin | P seen | PV issued | QK issued")
for j in range(8, 24):
    for g in range(2):
        a = t[g, j]
        print(f"j={j:2d} WG{g}: start {a[0]:7d} wait {a[1]-a[0]:5d} ld {a[2]-a[1]:4d} max {a[5]-a[2]:4d} turn {a[6]-a[5]:5d} exp {a[7]-a[6]:5d} pack {a[3]-a[7]:4d} store+arrive {a[4]-a[3]:4d} | period {t[g, j+1, 0]-a[0]:5d}")
    m = t[2, j]
    for g in range(2):
        print(f"      MMA{g}: start {m[g*4]:7d} waitP {m[g*4+1]-m[g*4]:5d} pv_issue {m[g*4+2]-m[g*4+1]:4d} qk_issue {m[g*4+3]-m[g*4+2]:4d}")
per = (t[0, 40, 0] - t[0, 8, 0]) / 32
w = np.mean(t[:2, 8:40, 1] - t[:2, 8:40, 0]); sm = np.mean(t[:2, 8:40, 3] - t[:2, 8:40, 2])
d = lambda a, b: np.mean(t[:2, 8:40, a] - t[:2, 8:40, b])
print(f"softmax split: max {d(5, 2):.0f}  wait-for-turn {d(6, 5):.0f}  exp {d(7, 6):.0f}  pack {d(3, 7):.0f}")
ld = np.mean(t[:2, 8:40, 2] - t[:2, 8:40, 1]); st = np.mean(t[:2, 8:40, 4] - t[:2, 8:40, 3])
print(f"mean period {per:.0f} cycles per step; per WG: wait {w:.0f}  ld {ld:.0f}  softmax {sm:.0f}  store+arrive {st:.0f}")
