"""Generate the golden fixtures under tests/golden/ (run ON THE B200 BOX, where the reference's own
CUDA sources - compiled to oracle/_ref/*.so by oracle/build_ref.py - and its Triton per-thread
quantizer (staged copy under the git-ignored oracle/_ref/stage/) can execute):

    python tests/golden/make_golden.py            # writes gpurun_out/golden/*.pt
    cp gpurun_out/golden/*.pt tests/golden/       # then commit

The reference ships no tests or known-answer vectors for this path, so these
outputs OF THE REFERENCE'S OWN CODE on seeded inputs are the pins: tests/test_oracle_golden.py
(CPU) checks oracle/sage_oracle.py against them, tests/test_golden_gpu.py checks the sm_100a
library against them bit-for-bit (codes/scales) or within the stated tolerance (attention output).
Every tensor is stored on CPU; inputs are stored too so nothing depends on RNG reproducibility.
"""
import importlib.util
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import build_ref  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out", "golden")
DEV = "cuda:0"


def load_triton_ref():
    return build_ref.load_triton("quant_per_thread")


def make_inputs(b, hq, hk, nq, nk, d, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(b, hq, nq, d, generator=g)
    k = torch.randn(b, hk, nk, d, generator=g) + 2.0 * torch.randn(1, hk, 1, d, generator=g)
    v = torch.randn(b, hk, nk, d, generator=g) + 0.5 * torch.randn(1, hk, 1, d, generator=g)
    return q.to(dtype), k.to(dtype), v.to(dtype)


def main():
    os.makedirs(OUT, exist_ok=True)
    rf = build_ref.load("fused")
    r80 = build_ref.load("sm80")
    r89 = build_ref.load("sm89")
    tri = load_triton_ref()
    print("refs:", rf is not None, r80 is not None, r89 is not None, tri is not None)
    cases = [
        ("a", 1, 2, 2, 200, 200, 128, torch.float16, False),
        ("b", 1, 2, 1, 130, 257, 64, torch.bfloat16, False),
        ("c", 1, 2, 2, 192, 192, 64, torch.float16, True),
        ("d", 1, 2, 2, 256, 256, 128, torch.bfloat16, True),
    ]
    for name, b, hq, hk, nq, nk, d, dtype, causal in cases:
        q, k, v = make_inputs(b, hq, hk, nq, nk, d, dtype, seed=ord(name))
        rec = dict(q=q, k=k, v=v, causal=causal, sm_scale=d ** -0.5)
        qg, kg, vg = q.to(DEV), k.to(DEV), v.to(DEV)
        km = kg.mean(dim=2)                     # core.py:612 (torch reduction, input dtype)
        rec["km"] = km.cpu()
        # ---- per_warp (quant.py:106-181 over the compiled launchers), WARPQ 32 and 16
        for wq in (32, 16):
            q8 = torch.empty_like(qg, dtype=torch.int8)
            k8 = torch.empty_like(kg, dtype=torch.int8)
            qs = torch.zeros((b, hq, (nq + 127) // 128 * (128 // wq)), device=DEV, dtype=torch.float32)
            ks = torch.zeros((b, hk, (nk + 63) // 64), device=DEV, dtype=torch.float32)
            rf.quant_per_warp_int8_cuda(qg, q8, qs, 128, wq, 1)
            rf.quant_per_block_int8_fuse_sub_mean_cuda(kg, km, k8, ks, 64, 1)
            rec[f"per_warp{wq}"] = tuple(t.cpu() for t in (q8, qs, k8, ks))
        # ---- per_block with sm_scale (quant.py:23-104)
        q8 = torch.empty_like(qg, dtype=torch.int8)
        k8 = torch.empty_like(kg, dtype=torch.int8)
        qs = torch.zeros((b, hq, (nq + 127) // 128), device=DEV, dtype=torch.float32)
        ks = torch.zeros((b, hk, (nk + 63) // 64), device=DEV, dtype=torch.float32)
        rf.quant_per_block_int8_sm_scale_cuda(qg, q8, qs, (d ** -0.5) * 1.44269504, 128, 1)
        rf.quant_per_block_int8_cuda(kg, k8, ks, 64, 1)
        rec["per_block"] = tuple(t.cpu() for t in (q8, qs, k8, ks))
        # ---- per_thread (Triton reference)
        if tri is not None:
            out = tri.per_thread_int8(qg, kg, km.unsqueeze(2))
            torch.cuda.synchronize()
            rec["per_thread"] = tuple(t.cpu() for t in out)
        # ---- V: sub_mean, per_channel_fp8 (with the caller's pad-to-128, core.py:822-828)
        vm = vg.mean(dim=2)
        vsm = torch.empty(vg.shape, dtype=torch.float16, device=DEV)
        rf.sub_mean_cuda(vg, vm, vsm, 1)
        rec["sub_mean"] = (vsm.cpu(), vm.cpu())
        n128 = (nk + 127) // 128 * 128
        vpad = torch.cat([vg, torch.zeros(b, hk, n128 - nk, d, dtype=dtype, device=DEV)], dim=2) if n128 > nk else vg
        for smooth, smax in ((False, 448.0), (True, 448.0), (False, 2.25)):
            vt = torch.zeros((b, hk, d, n128), dtype=dtype, device=DEV)
            rf.transpose_pad_permute_cuda(vpad, vt, 1)
            v8 = torch.zeros((b, hk, d, n128), dtype=torch.float8_e4m3fn, device=DEV)
            vs = torch.zeros((b, hk, d), dtype=torch.float32, device=DEV)
            vmean = torch.zeros((b, hk, d), dtype=torch.float32, device=DEV)
            if smooth:
                rf.mean_scale_fuse_quant_cuda(vt, v8, vmean, vs, n128, smax, 1)
            else:
                rf.scale_fuse_quant_cuda(vt, v8, vs, n128, smax, 1)
            rec[f"fp8_{int(smooth)}_{smax}"] = (v8.view(torch.uint8).cpu(), vs.cpu(), vmean.cpu() if smooth else None)
        # ---- attention: the reference's sm80 kernel (fp16 P.V) on per_warp32 and per_thread codes
        if r80 is not None and dtype == torch.float16:
            for gran, key in ((2, "per_warp32"), (3, "per_thread")):
                if key not in rec:
                    continue
                q8, qs, k8, ks = (t.to(DEV) for t in rec[key])
                o = torch.empty(qg.shape, dtype=dtype, device=DEV)
                lse = r80.qk_int8_sv_f16_accum_f32_attn(q8, k8, vg.to(torch.float16), o, qs, ks, 1, int(causal), gran,
                                                        d ** -0.5, 1)
                rec[f"sm80_f32_{key}"] = (o.cpu(), lse.cpu())
        if r80 is not None and dtype == torch.bfloat16:
            for gran, key in ((2, "per_warp32"), (3, "per_thread")):
                if key not in rec:
                    continue
                q8, qs, k8, ks = (t.to(DEV) for t in rec[key])
                o = torch.empty(qg.shape, dtype=dtype, device=DEV)
                lse = r80.qk_int8_sv_f16_accum_f32_attn(q8, k8, vg.to(torch.float16), o, qs, ks, 1, int(causal), gran,
                                                        d ** -0.5, 1)
                rec[f"sm80_f32_{key}"] = (o.cpu(), lse.cpu())
        # ---- attention: the reference's sm89 kernel (fp8 P.V, fused v_scale)
        if r89 is not None:
            v8u, vs, _ = rec["fp8_0_448.0"]
            v8 = v8u.to(DEV).view(torch.float8_e4m3fn)
            for gran, key in ((2, "per_warp32"), (3, "per_thread")):
                if key not in rec:
                    continue
                q8, qs, k8, ks = (t.to(DEV) for t in rec[key])
                o = torch.empty(qg.shape, dtype=dtype, device=DEV)
                lse = r89.qk_int8_sv_f8_accum_f32_fuse_v_scale_attn(q8, k8, v8, o, qs, ks, vs.to(DEV), 1, int(causal),
                                                                     gran, d ** -0.5, 1)
                rec[f"sm89_f32_{key}"] = (o.cpu(), lse.cpu())
        torch.cuda.synchronize()
        path = os.path.join(OUT, f"case_{name}.pt")
        torch.save(rec, path)
        print("wrote", path, os.path.getsize(path), "bytes", sorted(rec.keys()))


if __name__ == "__main__":
    main()
