"""Stray-write checks with guard regions (compute-sanitizer is closed on this GPU pool, so the new store paths of round 2
are checked directly): the output of a kernel sits in the middle of a larger buffer filled with a sentinel, and the
sentinel must be intact afterwards - in the guard rows before and after the tensor and, for the packed variable-length
kernel, in every row the kernel must not own.
"""
import pytest
import torch

import sageattention_b200 as sb
from sageattention_b200 import _lib, core
from sageattention_b200.quant import _dims, _dtype_code, _stream

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SENT = -12345.0


def guarded(shape, dtype, pad_rows=64):
    """A tensor of `shape` (rows = dim 0) inside a buffer with `pad_rows` sentinel rows on either side."""
    big = torch.full((shape[0] + 2 * pad_rows,) + tuple(shape[1:]), SENT, dtype=dtype, device=DEV)
    return big, big[pad_rows:pad_rows + shape[0]]


@pytest.mark.parametrize("d", [128, 64])
@pytest.mark.parametrize("causal", [False, True])
def test_varlen_attention_writes_only_its_own_rows(d, causal):
    """Tiles of 128 query rows are stored at arbitrary token offsets of the pack: whole tiles through TMA, the last
    tile of a sequence row by row.  Nothing outside [0, Tq) may be written, and every row inside must be."""
    lens = [130, 1, 257, 64, 300]           # partial last tiles of every size, a one-row sequence
    hq, hk = 4, 2
    tq = sum(lens)
    g = torch.Generator(device=DEV).manual_seed(d)
    q = torch.randn(tq, hq, d, generator=g, device=DEV).half()
    k = torch.randn(tq, hk, d, generator=g, device=DEV).half()
    v = torch.randn(tq, hk, d, generator=g, device=DEV).half()
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
    ref = sb.sageattn_varlen(q, k, v, cu, cu, max(lens), max(lens), is_causal=causal)
    km = sb.seq_mean(k.unsqueeze(0), tensor_layout="NHD")
    q8, qs, k8, ks, cqs, cks, (cq, ck, cp) = sb.per_block_int8_varlen(q, k, cu, cu, max(lens), max(lens), km=km)
    big, o = guarded((tq, hq, d), torch.float16)
    core._launch(False, q8.unsqueeze(0), k8.unsqueeze(0), v.unsqueeze(0), o.unsqueeze(0), qs, ks, None, None, "NHD", causal,
                 _lib.GRAN_PER_BLOCK, 32, 1.0 / 1.44269504088896340736, False, varlen=(cq, ck, cqs, cks, cp))
    torch.cuda.synchronize()
    assert (big[:64] == SENT).all() and (big[64 + tq:] == SENT).all(), "write outside the packed output"
    assert (o != SENT).all(), "a row of the pack was not written"
    assert torch.equal(o, ref)


def test_varlen_attention_leaves_rows_of_empty_query_sequences_alone():
    """A query row range the pack does not contain (cu_seqlens_q covering only part of the buffer) stays untouched."""
    lens_q, lens_k = [100, 0, 200], [128, 64, 200]
    tq, tk, h, d = 300, 392, 2, 128
    g = torch.Generator(device=DEV).manual_seed(3)
    q = torch.randn(tq + 50, h, d, generator=g, device=DEV).half()      # 50 tokens beyond the last sequence
    k = torch.randn(tk, h, d, generator=g, device=DEV).half()
    v = torch.randn(tk, h, d, generator=g, device=DEV).half()
    cq = torch.tensor([0, 100, 100, 300], dtype=torch.int32, device=DEV)
    ck = torch.tensor([0, 128, 192, 392], dtype=torch.int32, device=DEV)
    q8, qs, k8, ks, cqs, cks, (cq_, ck_, cp) = sb.per_block_int8_varlen(q, k, cq, ck, 200, 200)
    o = torch.full((tq + 50, h, d), SENT, dtype=torch.float16, device=DEV)
    core._launch(False, q8.unsqueeze(0), k8.unsqueeze(0), v.unsqueeze(0), o.unsqueeze(0), qs, ks, None, None, "NHD", False,
                 _lib.GRAN_PER_BLOCK, 32, 1.0 / 1.44269504088896340736, False, varlen=(cq_, ck_, cqs, cks, cp))
    torch.cuda.synchronize()
    assert (o[tq:] == SENT).all(), "rows beyond the last sequence were written"
    assert (o[:tq] != SENT).all()


@pytest.mark.parametrize("layout", ["HND", "NHD"])
@pytest.mark.parametrize("n,d", [(1000, 128), (130, 64), (77, 96)])
def test_v16_prologue_writes_only_the_f16_copy(layout, n, d):
    """sage_prepass_qk_int8_v16: the extra clusters that convert V write exactly v16's rows (ragged last tile)."""
    b, hq, hk = 2, 4, 2
    g = torch.Generator(device=DEV).manual_seed(n)
    mk = lambda h_: torch.randn(b, h_, n, d, generator=g, device=DEV).to(torch.bfloat16)
    q, k, v = mk(hq), mk(hk), mk(hk)
    if layout == "NHD":
        q, k, v = (t.transpose(1, 2).contiguous() for t in (q, k, v))
    fold = d not in (64, 128)
    ref = sb.prepass_qk_int8(q, k, "per_thread", tensor_layout=layout, _pad_head_dim=fold, v=v)
    # the same call through the C ABI with a guarded v16 buffer (flattened to rows of the last dimension)
    rows = v.numel() // d
    big, flat = guarded((rows, d), torch.float16, pad_rows=256)
    v16 = flat.view(v.shape)
    lib = _lib.load()
    bb, hq_, nq, d_og, q4 = _dims(q, layout)
    _, hk_, nk, _, k4 = _dims(k, layout)
    _, _, _, _, v4 = _dims(v, layout)
    _, _, _, _, vo4 = _dims(v16, layout)
    dd = 64 if d <= 64 else 128
    q8 = torch.empty(q.shape[:-1] + (dd,), dtype=torch.int8, device=DEV)
    k8 = torch.empty(k.shape[:-1] + (dd,), dtype=torch.int8, device=DEV)
    _, _, _, _, qo4 = _dims(q8, layout)
    _, _, _, _, ko4 = _dims(k8, layout)
    qpb, kpb = (nq + 127) // 128 * 32, (nk + 63) // 64 * 4
    qs = torch.empty(bb, hq_, qpb, device=DEV)
    ks = torch.empty(bb, hk_, kpb, device=DEV)
    km = torch.empty(bb, hk_, dd, dtype=k.dtype, device=DEV)
    _lib.check(lib.sage_prepass_qk_int8_v16(q4, k4, qo4, ko4, qs.data_ptr(), ks.data_ptr(), km.data_ptr(), v4, vo4, bb, hq_,
                                            hk_, nq, nk, dd, _dtype_code(q), 1, 32, qpb, kpb, d_og, _stream(q)))
    torch.cuda.synchronize()
    assert (big[:256] == SENT).all() and (big[256 + rows:] == SENT).all(), "write outside v16"
    assert torch.equal(v16, ref[5]) and torch.equal(q8, ref[0]) and torch.equal(k8, ref[2])
