"""development: isolate the o_peer_* epilogue (run on a GPU box): python tools/dbg_opeers.py ROWS D"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sageattention_b200 import parallel, quant, core, _lib
rows, d = int(sys.argv[1]), int(sys.argv[2])
n, h = 1000, 2
g = torch.Generator().manual_seed(0)
q, k, v = (torch.randn(1, n, h, d, generator=g).to(torch.bfloat16).cuda() for _ in range(3))
km = quant.seq_mean(k, "NHD")
q8, qs, k8, ks = quant.per_thread_int8(q, k, km, tensor_layout="NHD")
v8, vs, _ = quant.per_channel_fp8(v, tensor_layout="NHD", scale_max=448.0, smooth_v=False, _n_stat=1024)
o_ref = torch.empty((1, n, h, d), dtype=q.dtype, device="cuda")
core._launch(True, q8, k8, v8, o_ref, qs, ks, vs, None, "NHD", False, _lib.GRAN_PER_THREAD, 32, d ** -0.5, False)
torch.cuda.synchronize(); print("plain ok", flush=True)
nd = (n + rows - 1) // rows
dests = [torch.full((rows, h, d), 7.0, dtype=q.dtype, device="cuda") for _ in range(nd)]
scratch = torch.empty(16 * 136, dtype=torch.uint8, device="cuda")
o = torch.empty((1, 128, h, d), dtype=q.dtype, device="cuda")
core._launch(True, q8, k8, v8, o, qs, ks, vs, None, "NHD", False, _lib.GRAN_PER_THREAD, 32, d ** -0.5, False,
             o_peers=([t.data_ptr() for t in dests], rows, d, h * d, scratch))
torch.cuda.synchronize(); print("peers launched ok", flush=True)
got = torch.cat(dests)
print("equal", torch.equal(got[:n], o_ref[0]), "tail untouched", bool((got[n:] == 7.0).all()))
