import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _load_ref(which):
    from oracle import build_ref
    try:
        return build_ref.load(which)
    except Exception as e:  # pragma: no cover - diagnostic only
        print(f"[conftest] could not load reference oracle {which}: {e}")
        return None


@pytest.fixture(scope="session")
def ref_fused():
    m = _load_ref("fused")
    if m is None:
        pytest.skip("oracle/_ref/_ref_fused.so not built (python oracle/build_ref.py fused)")
    return m


@pytest.fixture(scope="session")
def ref_sm80():
    m = _load_ref("sm80")
    if m is None:
        pytest.skip("oracle/_ref/_ref_sm80.so not built")
    return m


@pytest.fixture(scope="session")
def ref_sm89():
    m = _load_ref("sm89")
    if m is None:
        pytest.skip("oracle/_ref/_ref_sm89.so not built")
    return m


@pytest.fixture(scope="session")
def ref_triton():
    """Loader of the reference's staged Triton modules (oracle/_ref/stage, see oracle/build_ref.py)."""
    from oracle import build_ref

    def get(stem):
        m = build_ref.load_triton(stem)
        if m is None:
            pytest.skip(f"oracle/_ref/stage/{stem}.py not staged (python oracle/build_ref.py triton)")
        return m
    return get


def accuracy(a, b):
    """(cosine similarity, relative L1, rmse) of a against b."""
    import torch
    a = a.float().flatten()
    b = b.float().flatten()
    cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
    l1 = ((a - b).abs().sum() / b.abs().sum()).item()
    rmse = torch.sqrt(torch.mean((a - b) ** 2)).item()
    return cos, l1, rmse


def exact_attention_from_codes(q8, k8, v, q_scale, k_scale, gran, causal, sm_scale, v_scale=None, warp_q=32):
    """fp64 attention computed from the SAME int8 codes / scales (and the dequantised V the kernel
    sees): the ground truth both a kernel under test and the reference kernel are rounding.
    v: f16 [B,Hkv,Nk,D], or e4m3 [B,Hkv,D,Npad] in the reference's permuted layout (+ v_scale)."""
    import torch
    from oracle import sage_oracle as O
    b, hq, nq, d = q8.shape
    _, hk, nk, _ = k8.shape
    g = hq // hk
    dev = q8.device
    qi = O._scale_index(nq, gran, True, warp_q).clamp_max(q_scale.size(-1) - 1).to(dev)
    ki = O._scale_index(nk, gran, False).clamp_max(k_scale.size(-1) - 1).to(dev)
    s = torch.matmul(q8.double(), k8.double().repeat_interleave(g, 1).transpose(2, 3))
    s = s * q_scale.double()[:, :, qi][..., None] * k_scale.double()[:, :, ki].repeat_interleave(g, 1)[:, :, None, :]
    s = s * sm_scale
    if causal:
        s = s.masked_fill(torch.triu(torch.ones(nq, nk, dtype=torch.bool, device=dev), 1), float("-inf"))
    p = torch.softmax(s, dim=-1)
    if v.dtype == torch.float8_e4m3fn:
        inv = torch.tensor(O.PERM16).argsort().to(dev)
        npad = v.size(-1)
        vt = v.float().reshape(b, hk, d, npad // 16, 16)[..., inv].reshape(b, hk, d, npad)
        vf = vt.transpose(2, 3)[:, :, :nk].double() * v_scale.double()[:, :, None, :]
    else:
        vf = v.double()
    return torch.matmul(p, vf.repeat_interleave(g, 1)), torch.logsumexp(s, -1) * 1.4426950408889634
