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
