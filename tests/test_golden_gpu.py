"""GPU: the sm_100a library (through the C ABI) against the golden fixtures - outputs of the
reference's own compiled kernels (tests/golden/make_golden.py).  Codes and scales: BIT-EXACT.
Attention outputs (floating point, different tile order and f32 accumulation in TMEM):
cos >= 0.9999, rel-L1 <= 5e-3 (fp16 PV vs the sm80 kernel); fp8 PV vs the sm89 kernel: both round P
to e4m3 against different reference maxima (independent rounding errors), so cos >= 0.9995,
rel-L1 <= 3e-2 pairwise AND our distance to the fp64 result from the same codes must be within
1.10x of the reference kernel's own; LSE within 2e-3."""
import glob
import os

import pytest
import torch

import sageattention_b200 as sb
from sageattention_b200 import core
from conftest import accuracy, exact_attention_from_codes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(glob.glob(os.path.join(HERE, "golden", "case_*.pt")))


def _load(path):
    return torch.load(path, map_location="cpu", weights_only=False)


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_quantizers_bit_exact_vs_golden(path):
    g = _load(path)
    q, k, v, km = (g[x].to(DEV) for x in ("q", "k", "v", "km"))
    nq = q.size(2)
    for wq in (32, 16):
        r = g[f"per_warp{wq}"]
        o = sb.per_warp_int8(q, k, km, WARPQ=wq)
        nw = (nq + wq - 1) // wq
        assert torch.equal(o[0].cpu(), r[0]) and torch.equal(o[2].cpu(), r[2])
        assert torch.equal(o[1].cpu()[..., :nw], r[1][..., :nw]) and torch.equal(o[3].cpu(), r[3])
    r = g["per_block"]
    o = sb.per_block_int8(q, k, None, sm_scale=g["sm_scale"])
    for a, b in zip(o, r):
        assert torch.equal(a.cpu(), b)
    r = g["per_thread"]
    o = sb.per_thread_int8(q, k, km)
    for a, b in zip(o, r):
        assert torch.equal(a.cpu(), b)
    assert torch.equal(sb.seq_mean(k).cpu(), g["km"]), "seq_mean vs torch mean"
    sm, _ = sb.sub_mean(v, vm=g["sub_mean"][1].to(DEV))
    assert torch.equal(sm.cpu(), g["sub_mean"][0])
    n128 = (k.size(2) + 127) // 128 * 128
    for smooth, smax in ((False, 448.0), (True, 448.0), (False, 2.25)):
        r8, rs, rm = g[f"fp8_{int(smooth)}_{smax}"]
        o8, os_, om = sb.per_channel_fp8(v, scale_max=smax, smooth_v=smooth, _n_stat=n128)
        assert torch.equal(o8.view(torch.uint8).cpu(), r8), "fp8 codes"
        assert torch.equal(os_.cpu(), rs), "v_scale"
        if smooth:
            assert torch.equal(om.cpu(), rm), "v_mean"


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_attention_vs_golden_reference_kernels(path):
    g = _load(path)
    dtype = g["q"].dtype
    causal, sm = g["causal"], g["sm_scale"]
    for gran, key in ((2, "per_warp32"), (3, "per_thread")):
        q8, qs, k8, ks = (t.to(DEV) for t in g[key])
        o = torch.empty(q8.shape, dtype=dtype, device=DEV)
        lse = core._launch(False, q8, k8, g["v"].to(DEV).to(torch.float16), o, qs, ks, None, None, "HND", causal,
                           gran, 32, sm, True)
        ro, rlse = g[f"sm80_f32_{key}"]
        cos, l1, _ = accuracy(o.cpu(), ro)
        assert cos >= 0.9999 and l1 <= 5e-3, (key, cos, l1)
        assert (lse.cpu() - rlse).abs().max().item() <= 2e-3
        v8u, vs, _ = g["fp8_0_448.0"]
        lse = core._launch(True, q8, k8, v8u.to(DEV).view(torch.float8_e4m3fn), o, qs, ks, vs.to(DEV), None, "HND",
                           causal, gran, 32, sm, True)
        ro, rlse = g[f"sm89_f32_{key}"]
        cos, l1, _ = accuracy(o.cpu(), ro)
        assert cos >= 0.9995 and l1 <= 3e-2, (key, cos, l1)
        assert (lse.cpu() - rlse).abs().max().item() <= 2e-3
        exact, _ = exact_attention_from_codes(q8, k8, v8u.to(DEV).view(torch.float8_e4m3fn), qs, ks, gran, causal,
                                              sm, v_scale=vs.to(DEV))
        _, l1_mine, _ = accuracy(o, exact)
        _, l1_ref, _ = accuracy(ro.to(DEV), exact)
        assert l1_mine <= 1.10 * l1_ref + 1e-4, (key, l1_mine, l1_ref)
