"""First-contact diagnostics on the B200 box (not a test: prints, never asserts).
Each attention case runs in its own subprocess so a trap / hang in one does not hide the others.
Usage: python tools/gpu_probe.py [quant|attn|case <json>] """
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def acc(a, b):
    import torch
    a = a.float().flatten(); b = b.float().flatten()
    return (torch.nn.functional.cosine_similarity(a, b, dim=0).item(),
            ((a - b).abs().sum() / b.abs().sum().clamp_min(1e-20)).item())


def quant():
    import torch
    import sageattention_b200 as sb
    from oracle import build_ref, sage_oracle as O
    rf = build_ref.load("fused")
    dev = "cuda:0"
    print("ref fused:", rf is not None)
    for dtype in (torch.float16, torch.bfloat16):
        for (n, d) in ((1000, 128), (333, 64)):
            g = torch.Generator().manual_seed(n)
            q = (torch.randn(2, 3, n, d, generator=g) + 2 * torch.randn(1, 3, 1, d, generator=g)).to(dtype).to(dev)
            k = (torch.randn(2, 3, n, d, generator=g) + 2 * torch.randn(1, 3, 1, d, generator=g)).to(dtype).to(dev)
            km = k.mean(dim=2)
            mine = sb.per_warp_int8(q, k, km)
            torch.cuda.synchronize()
            if rf is not None:
                q8 = torch.empty_like(q, dtype=torch.int8); k8 = torch.empty_like(k, dtype=torch.int8)
                qs = torch.zeros_like(mine[1]); ks = torch.zeros_like(mine[3])
                rf.quant_per_warp_int8_cuda(q, q8, qs, 128, 32, 1)
                rf.quant_per_block_int8_fuse_sub_mean_cuda(k, km, k8, ks, 64, 1)
                nw = (n + 31) // 32
                print(dtype, n, d, "per_warp vs ref: q codes neq", (mine[0] != q8).sum().item(), "q scale neq",
                      (mine[1][..., :nw] != qs[..., :nw]).sum().item(), "k codes neq", (mine[2] != k8).sum().item(),
                      "k scale neq", (mine[3] != ks).sum().item())
            o = O.per_warp_int8(q.cpu(), k.cpu(), km.cpu())
            print("   per_warp vs oracle: q neq", (mine[0].cpu() != o[0]).sum().item(), "k neq", (mine[2].cpu() != o[2]).sum().item(),
                  "qs neq", (mine[1].cpu() != o[1]).sum().item(), "ks neq", (mine[3].cpu() != o[3]).sum().item())
            mt = sb.per_thread_int8(q, k, km)
            ot = O.per_thread_int8(q.cpu(), k.cpu(), km.cpu())
            print("   per_thread(div=%s) vs oracle: q neq" % os.environ.get("SAGE_PT_DIV", "1"), (mt[0].cpu() != ot[0]).sum().item(),
                  "k neq", (mt[2].cpu() != ot[2]).sum().item(), "qs neq", (mt[1].cpu() != ot[1]).sum().item(),
                  "ks neq", (mt[3].cpu() != ot[3]).sum().item())
            stage = os.path.join(ROOT, "oracle", "_ref", "stage", "quant_per_thread.py")
            if os.path.exists(stage):
                import importlib.util
                spec = importlib.util.spec_from_file_location("ref_qpt", stage)
                m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
                rt = m.per_thread_int8(q, k, km.unsqueeze(2))
                torch.cuda.synchronize()
                print("   per_thread vs TRITON ref: q neq", (mt[0] != rt[0]).sum().item(), "qs neq", (mt[1] != rt[1]).sum().item(),
                      "k neq", (mt[2] != rt[2]).sum().item(), "ks neq", (mt[3] != rt[3]).sum().item())
            for smooth in (False, True):
                v8, vs, vm = sb.per_channel_fp8(k, smooth_v=smooth)
                torch.cuda.synchronize()
                if rf is not None:
                    npad = (n + 63) // 64 * 64
                    vt = torch.zeros((2, 3, d, npad), dtype=dtype, device=dev)
                    rf.transpose_pad_permute_cuda(k, vt, 1)
                    r8 = torch.zeros((2, 3, d, npad), dtype=torch.float8_e4m3fn, device=dev)
                    rs = torch.zeros((2, 3, d), dtype=torch.float32, device=dev)
                    rm = torch.zeros((2, 3, d), dtype=torch.float32, device=dev)
                    if smooth:
                        rf.mean_scale_fuse_quant_cuda(vt, r8, rm, rs, n, 448.0, 1)
                    else:
                        rf.scale_fuse_quant_cuda(vt, r8, rs, n, 448.0, 1)
                    print("   fp8 smooth=%s vs ref: codes neq" % smooth, (v8.view(torch.uint8) != r8.view(torch.uint8)).sum().item(),
                          "scale neq", (vs != rs).sum().item(), "mean neq", (vm != rm).sum().item() if smooth else "-")
                o8, os_, om = O.per_channel_fp8(k.cpu(), smooth_v=smooth)
                print("   fp8 smooth=%s vs oracle: codes neq" % smooth, (v8.cpu().view(torch.uint8) != o8.view(torch.uint8)).sum().item(),
                      "of", o8.numel(), "scale neq", (vs.cpu() != os_).sum().item(), "mean neq", (vm.cpu() != om).sum().item() if smooth else "-")
            sm = sb.seq_mean(k)
            print("   seq_mean vs torch neq", (sm != km).sum().item(), "of", km.numel(), "maxdiff", (sm.float() - km.float()).abs().max().item())


def case(cfg):
    import torch
    import sageattention_b200 as sb
    from sageattention_b200 import core
    from oracle import sage_oracle as O
    dev = "cuda:0"
    b, hq, hk, nq, nk, d, causal, pv, gran = (cfg[k] for k in ("b", "hq", "hk", "nq", "nk", "d", "causal", "pv", "gran"))
    dtype = torch.float16
    g = torch.Generator().manual_seed(1)
    q = torch.randn(b, hq, nq, d, generator=g).to(dtype).to(dev)
    k = (torch.randn(b, hk, nk, d, generator=g) + 2 * torch.randn(1, hk, 1, d, generator=g)).to(dtype).to(dev)
    v = torch.randn(b, hk, nk, d, generator=g).to(dtype).to(dev)
    km = k.mean(dim=2)
    q8, qs, k8, ks = (sb.per_thread_int8 if gran == 3 else sb.per_warp_int8)(q, k, km)
    if pv == "f8":
        vq, vs, _ = sb.per_channel_fp8(v, smooth_v=False, _n_stat=(nk + 127) // 128 * 128)
    else:
        vq, vs = v, None
    o = torch.zeros(q8.shape, dtype=dtype, device=dev)
    nqp = (nq + 255) // 256 * 256
    dbg = torch.full((b, hq, nqp, 128), -777, dtype=torch.int32, device=dev)
    lse = core._launch(pv == "f8", q8, k8, vq, o, qs, ks, vs, None, "HND", causal, gran, 32, d ** -0.5, True, debug_s=dbg)
    torch.cuda.synchronize()
    # S check (kv tile 0)
    kk = min(nk, 128)
    sref = torch.matmul(q8.float(), k8[:, :, :kk].float().repeat_interleave(hq // hk, 1).transpose(2, 3)).int()
    sgot = dbg[:, :, :nq, :kk]
    print("  S tile0 mismatches:", (sgot != sref).sum().item(), "of", sref.numel())
    if (sgot != sref).any():
        print("   S got[0,0,:4,:8]\n", sgot[0, 0, :4, :8].cpu(), "\n   S ref\n", sref[0, 0, :4, :8].cpu())
    oo, olse = O.attn_int8(q8.cpu(), k8.cpu(), vq.cpu(), qs.cpu(), ks.cpu(), pv=pv, v_scale=None if vs is None else vs.cpu(),
                           gran=gran, is_causal=causal, sm_scale=d ** -0.5, return_lse=True, out_dtype=dtype)
    print("  O vs oracle cos/l1:", acc(o.cpu(), oo), " lse maxdiff", (lse.cpu() - olse).abs().max().item(),
          " finite", torch.isfinite(o).all().item())
    ref = torch.nn.functional.scaled_dot_product_attention(q.float(), k.float().repeat_interleave(hq // hk, 1),
                                                           v.float().repeat_interleave(hq // hk, 1), is_causal=causal)
    print("  O vs sdpa   cos/l1:", acc(o, ref), " oracle vs sdpa", acc(oo, ref.cpu()))
    if acc(o.cpu(), oo)[0] < 0.999:
        print("   o[0,0,:2,:8]", o[0, 0, :2, :8].cpu(), "\n   oracle", oo[0, 0, :2, :8])


def attn():
    cases = []
    for pv in ("f8", "f16"):
        for d in (128, 64):
            cases.append(dict(b=1, hq=1, hk=1, nq=128, nk=128, d=d, causal=False, pv=pv, gran=2))
            cases.append(dict(b=1, hq=2, hk=1, nq=512, nk=512, d=d, causal=False, pv=pv, gran=3))
            cases.append(dict(b=2, hq=2, hk=2, nq=700, nk=900, d=d, causal=False, pv=pv, gran=3))
            cases.append(dict(b=1, hq=2, hk=2, nq=1024, nk=1024, d=d, causal=True, pv=pv, gran=3))
    for c in cases:
        print("CASE", c, flush=True)
        try:
            p = subprocess.run([sys.executable, __file__, "case", json.dumps(c)], timeout=120,
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            print(p.stdout[-3000:], "rc", p.returncode, flush=True)
        except subprocess.TimeoutExpired as e:
            print("  TIMEOUT", (e.stdout or b"")[-2000:], flush=True)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what == "case":
        case(json.loads(sys.argv[2]))
    elif what == "quant":
        quant()
    elif what == "attn":
        attn()
    else:
        for m in ("0", "1", "2") if False else ("1",):
            pass
        subprocess.run([sys.executable, __file__, "quant"])
        attn()
