"""CPU (`-m "not gpu"`), world_size 2 over gloo: the Ulysses all-to-all wrapper and head sharding
reproduce single-process attention exactly (stand-in attention = fp32 SDPA on CPU tensors; the
quantised operator itself needs a GPU and is covered by tests/test_parallel_gpu.py)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sageattention_b200 import parallel


def _sdpa_nhd(q, k, v, tensor_layout="NHD", is_causal=False, **kw):
    assert tensor_layout == "NHD"
    g = q.size(2) // k.size(2)
    o = torch.nn.functional.scaled_dot_product_attention(
        q.transpose(1, 2), k.transpose(1, 2).repeat_interleave(g, 1), v.transpose(1, 2).repeat_interleave(g, 1),
        is_causal=is_causal)
    return o.transpose(1, 2)


def _worker(rank, world, port, hq, hk, q_out, b=2):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        n, d = 24, 16
        q = torch.randn(b, n, hq, d, generator=g)
        k = torch.randn(b, n, hk, d, generator=g)
        v = torch.randn(b, n, hk, d, generator=g)
        full = _sdpa_nhd(q, k, v, is_causal=True)
        nl = n // world
        sl = slice(rank * nl, (rank + 1) * nl)
        o = parallel.ulysses_sageattn(q[:, sl].contiguous(), k[:, sl].contiguous(), v[:, sl].contiguous(),
                                      attn_fn=_sdpa_nhd, is_causal=True)
        err = (o - full[:, sl]).abs().max().item()
        # head sharding: no communication, concatenating rank outputs gives the full result
        qs, ks, vs = (parallel.shard_heads(t, rank, world, "NHD") for t in (q, k, v))
        oh = _sdpa_nhd(qs, ks, vs, is_causal=True)
        err2 = (oh - parallel.shard_heads(full, rank, world, "NHD")).abs().max().item()
        q_out.put((rank, err, err2, tuple(o.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("hq,hk,b", [(4, 4, 2), (8, 2, 2), (4, 4, 1), (8, 2, 1)])   # b = 1: the one-copy fast path
def test_ulysses_and_head_sharding_world2(hq, hk, b):
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + hq + 16 * b
    procs = [ctx.Process(target=_worker, args=(r, 2, port, hq, hk, q_out, b)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, err2, shape in res:
        assert shape == (b, 12, hq, 16)
        assert err < 1e-5, (rank, err)
        assert err2 < 1e-5, (rank, err2)


def test_world1_passthrough_and_errors():
    q = torch.randn(1, 8, 4, 16)
    o = parallel.ulysses_sageattn(q, q, q, attn_fn=_sdpa_nhd)
    assert torch.allclose(o, _sdpa_nhd(q, q, q))
    with pytest.raises(ValueError):
        parallel.ulysses_sageattn(q, q, q, attn_fn=_sdpa_nhd, return_lse=True)
    with pytest.raises(ValueError):
        parallel.shard_heads(q, 0, 3, "NHD")


# ---------------------------------------------------------------------------------------------------
# Quantised-payload Ulysses (payload="int8"): the host logic - ownership re-cut on 128-row boundaries, the halo from
# the next rank, the global K mean / V range, the send layout, the two inbound collectives, the re-assembly of scales and
# V^T by source rank - over gloo with a CPU stand-in for the device half, built from the oracle's quantisers and
# attention (test infrastructure).  The result must EQUAL the same stand-in pipeline run on the whole sequence by one
# process: any slip in the bookkeeping moves codes or scales to the wrong rows.
class _OracleOps:
    scale_max = 448.0

    def __init__(self):
        import sys
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from oracle import sage_oracle
        self.o = sage_oracle

    def stats(self, k, v):
        return torch.stack((k[0].float().sum(0), v[0].float().abs().amax(0)))

    def mean_from_sum(self, ksum, n_total, dtype):
        return (ksum * (1.0 / n_total)).to(dtype)

    def quant_rows(self, x, rows, km, is_key, world, out, scale):
        m, h, d = x.shape
        hp = h // world
        xb = x[:rows].permute(1, 0, 2).unsqueeze(0)            # [1, H, rows, D]
        q8, qs, k8, ks = self.o.per_thread_int8(xb, xb, km=km.reshape(1, h, d) if km is not None else None)
        c, s = (k8, ks) if is_key else (q8, qs)
        out[:, :rows].copy_(c[0].reshape(world, hp, rows, d).permute(0, 2, 1, 3))
        scale.copy_(s[0].reshape(world, hp, -1))

    def v_convert(self, x, rows, absmax, world, out, v_scale):
        m, h, d = x.shape
        hp = h // world
        n_stat = (rows + 127) // 128 * 128
        xb = torch.zeros((1, h, n_stat, d), dtype=x.dtype)
        xb[0, :, :rows] = x[:rows].permute(1, 0, 2)
        vt = self.o.transpose_pad_permute(xb, "HND").float()    # [1, H, D, n_stat]
        y = vt * (self.scale_max / absmax)[None, :, :, None]
        v8 = self.o._f32_to_e4m3_satfinite(y).view(torch.uint8)[0]
        out[..., :n_stat].copy_(v8.reshape(world, hp, d, n_stat))
        v_scale.copy_(absmax / self.scale_max)

    def attention(self, q8, k8, v8, q_scale, k_scale, v_scale, out_dtype, is_causal, sm_scale):
        n = q8.size(1)
        npad = (n + 127) // 128 * 128
        o = self.o.attn_int8(q8.permute(0, 2, 1, 3), k8.permute(0, 2, 1, 3),
                             v8[None, :, :, :npad].contiguous().view(torch.float8_e4m3fn), q_scale, k_scale, pv="f8",
                             v_scale=v_scale[None], is_causal=is_causal, sm_scale=sm_scale, out_dtype=out_dtype)
        return o.permute(0, 2, 1, 3).contiguous()


def _q8_worker(rank, world, port, nl, h, d, q_out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(1)
        n = nl * world
        q = torch.randn(1, n, h, d, generator=g).to(torch.float16)
        k = (torch.randn(1, n, h, d, generator=g) + 0.5).to(torch.float16)
        v = torch.randn(1, n, h, d, generator=g).to(torch.float16)
        ops = _OracleOps()
        # one process, whole sequence, same stand-in arithmetic (mean folded shard by shard like the ranks do)
        ksum = sum(k[0, r * nl:(r + 1) * nl].float().sum(0) for r in range(world))
        km = ops.mean_from_sum(ksum, n, k.dtype)
        vmax = v[0].float().abs().amax(0)
        q8 = torch.empty((1, n, h, d), dtype=torch.int8)
        k8 = torch.empty((1, n, h, d), dtype=torch.int8)
        tq, tk = parallel._scale_counts(n)
        qs = torch.empty((1, h, tq)); ks = torch.empty((1, h, tk))
        ops.quant_rows(q[0], n, None, False, 1, q8, qs)
        ops.quant_rows(k[0], n, km, True, 1, k8, ks)
        npad = (n + 127) // 128 * 128
        v8 = torch.zeros((1, h, d, npad), dtype=torch.uint8)
        vsc = torch.empty((h, d))
        ops.v_convert(v[0], n, vmax, 1, v8, vsc)
        full = ops.attention(q8, k8, v8[0], qs, ks, vsc, q.dtype, True, d ** -0.5)
        sl = slice(rank * nl, (rank + 1) * nl)
        o = parallel.ulysses_sageattn(q[:, sl].contiguous(), k[:, sl].contiguous(), v[:, sl].contiguous(),
                                      payload="int8", _q8_ops=ops, is_causal=True)
        q_out.put((rank, tuple(o.shape), bool(torch.equal(o, full[:, sl])), (o.float() - full[:, sl].float()).abs().max().item()))
    finally:
        dist.destroy_process_group()


def _q8_fallback_worker(rank, world, port, q_out):
    """Shards shorter than the halo they would need: payload="int8" must fall back to the 16-bit exchange."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(2)
        n, h, d = 80, 4, 64
        q, k, v = (torch.randn(1, n, h, d, generator=g) for _ in range(3))
        full = _sdpa_nhd(q, k, v)
        nl = n // world
        sl = slice(rank * nl, (rank + 1) * nl)
        assert parallel.q8_geometry(nl, world) is None
        o = parallel.ulysses_sageattn(q[:, sl].contiguous(), k[:, sl].contiguous(), v[:, sl].contiguous(),
                                      attn_fn=_sdpa_nhd, payload="int8", _q8_ops=_OracleOps())
        q_out.put((rank, (o - full[:, sl]).abs().max().item()))
    finally:
        dist.destroy_process_group()


def test_quantised_payload_falls_back_for_short_shards():
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = 33500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_q8_fallback_worker, args=(r, 2, port, q_out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err in res:
        assert err < 1e-5, (rank, err)


@pytest.mark.parametrize("world,nl", [(2, 300), (2, 256), (3, 520), (4, 500)])   # ragged (halo), aligned, 3 and 4 ranks
def test_ulysses_quantised_payload_host_logic(world, nl):
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = 31500 + (os.getpid() % 2000) + world * 7 + nl % 13
    h, d = 2 * world, 64
    procs = [ctx.Process(target=_q8_worker, args=(r, world, port, nl, h, d, q_out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, shape, same, err in res:
        assert shape == (1, nl, h, d)
        assert same, (rank, err)


def test_q8_geometry():
    m, own = parallel.q8_geometry(9450, 8)          # Wan2.1 on 8 GPUs
    assert m == 9472 and own == [9472] * 7 + [75600 - 7 * 9472] and sum(own) == 75600
    m, own = parallel.q8_geometry(14882, 8)         # HunyuanVideo on 8 GPUs
    assert m == 14976 and sum(own) == 119056 and all(o % 128 == 0 for o in own[:-1])
    assert parallel.q8_geometry(4096, 4) == (4096, [4096] * 4)
    assert parallel.q8_geometry(200, 4) is None     # shards shorter than the halo they would need
