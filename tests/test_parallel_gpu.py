"""GPU, 2 ranks over NCCL (skipped on a 1-GPU box): Ulysses sageattn on sequence-sharded inputs
equals the single-GPU operator on the gathered inputs BIT FOR BIT (same heads, same full sequence,
same kernels), and is within the operator's tolerance of fp32 SDPA."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q_out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import sageattention_b200 as sb
        from sageattention_b200 import parallel
        g = torch.Generator().manual_seed(0)
        b, n, h, d = 1, 1024, 8, 128
        q, k, v = (torch.randn(b, n, h, d, generator=g).to(torch.bfloat16).cuda() for _ in range(3))
        full = sb.sageattn(q, k, v, tensor_layout="NHD")
        nl = n // world
        sl = slice(rank * nl, (rank + 1) * nl)
        same = True
        # head groups pipelined against the transfers, over NCCL or over peer memory: same bits every way,
        # also when the same arena is used again by a second call
        for transport, chunks in (("nccl", 1), ("nccl", 2), ("nccl", 4), ("p2p", 1), ("p2p", 2), ("p2p", 4), ("p2p", 2)):
            o = parallel.ulysses_sageattn(q[:, sl].contiguous(), k[:, sl].contiguous(), v[:, sl].contiguous(),
                                          head_chunks=chunks, transport=transport)
            torch.cuda.synchronize()
            same = same and torch.equal(o, full[:, sl])
        ref = torch.nn.functional.scaled_dot_product_attention(q.float().transpose(1, 2), k.float().transpose(1, 2),
                                                               v.float().transpose(1, 2)).transpose(1, 2)[:, sl]
        cos = torch.nn.functional.cosine_similarity(o.float().flatten(), ref.flatten(), dim=0).item()
        q_out.put((rank, same, cos))
    finally:
        dist.destroy_process_group()


def test_ulysses_two_gpus_matches_single_gpu():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = 29700 + os.getpid() % 1000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q_out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, same, cos in res:
        assert same, f"rank {rank}: Ulysses output differs from the single-GPU operator"
        assert cos >= 0.999


def test_two_devices_from_one_process():
    """One process driving two GPUs (not the recommended layout, but the reference's core.py:583 supports it with
    torch.cuda.set_device): per-device state of the library (function attributes, scheduler counters, SM count)
    must not leak from the first device to the second."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import sageattention_b200 as sb
    g = torch.Generator().manual_seed(1)
    q, k, v = (torch.randn(1, 4, 700, 128, generator=g).to(torch.bfloat16) for _ in range(3))
    outs = []
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        o8 = sb.sageattn(q.to(dev), k.to(dev), v.to(dev), is_causal=True)
        o16 = sb.sageattn_qk_int8_pv_fp16_cuda(q.to(dev), k.to(dev), v.to(dev))
        torch.cuda.synchronize(dev)
        outs.append((o8.cpu(), o16.cpu()))
    for o8, o16 in outs[1:]:
        assert torch.equal(o8, outs[0][0]) and torch.equal(o16, outs[0][1])


# ---------------------------------------------------------------------------------------------------
# Quantised-payload Ulysses.  The device half runs on ONE GPU (these tests are not skipped on the
# driver's 1-GPU box): the C-ABI entry points that the path adds, and the existing quantisers driven through the
# "batch = destination rank" strides, against the one-GPU quantisers.  The end-to-end 2-rank test follows.
def _q8_inputs(n, h, d, dtype=torch.bfloat16, seed=3):
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(1, n, h, d, generator=g).to(dtype).cuda()
    k = (torch.randn(1, n, h, d, generator=g) + torch.randn(1, 1, h, d, generator=g)).to(dtype).cuda()
    v = (torch.randn(1, n, h, d, generator=g) * (1 + torch.rand(1, 1, h, d, generator=g))).to(dtype).cuda()
    return q, k, v


@pytest.mark.parametrize("d", [64, 128])
@pytest.mark.parametrize("n", [1000, 1152])
def test_q8_device_half_matches_one_gpu_quantisers(n, d):
    from sageattention_b200 import parallel, quant
    ops = parallel._Q8CudaOps()
    h, world = 6, 3
    hp = h // world
    q, k, v = _q8_inputs(n, h, d)
    # statistics: column sums / channel range of a shard
    st = ops.stats(k, v)
    ref_sum = k[0].double().sum(0)
    assert torch.allclose(st[0].double(), ref_sum, rtol=1e-5, atol=1e-3)
    assert torch.equal(st[1], v[0].float().abs().amax(0))
    km = ops.mean_from_sum(st[0], n, k.dtype)
    # the one-GPU mean (torch arithmetic) may differ from the folded sum in the last bit of a channel, no more
    km1 = quant.seq_mean(k, "NHD")[0]
    assert (km.float() - km1.float()).abs().max() <= 2 * torch.finfo(k.dtype).eps * km1.float().abs().max()
    # codes and scales, written through "batch = destination rank" strides into a [P, M, 2, hp, D] send buffer
    m = (n + 127) // 128 * 128
    send = torch.zeros((world, m, 2, hp, d), dtype=torch.int8, device="cuda")
    nsq, nsk = parallel._scale_counts(n)
    qs = torch.empty((world, hp, nsq), device="cuda")
    ks = torch.empty((world, hp, nsk), device="cuda")
    stage = torch.zeros((3, m, h, d), dtype=q.dtype, device="cuda")
    for i, x in enumerate((q, k, v)):
        stage[i, :n] = x[0]
    ops.quant_rows(stage[0], n, None, False, world, send[:, :, 0], qs)
    ops.quant_rows(stage[1], n, km, True, world, send[:, :, 1], ks)
    q8, q_scale, k8, k_scale = quant.per_thread_int8(q, k, km.reshape(1, 1, h, d), tensor_layout="NHD")
    assert torch.equal(send[:, :n, 0].permute(1, 0, 2, 3).reshape(n, h, d), q8[0])
    assert torch.equal(send[:, :n, 1].permute(1, 0, 2, 3).reshape(n, h, d), k8[0])
    assert torch.equal(qs.reshape(h, nsq), q_scale[0]) and torch.equal(ks.reshape(h, nsk), k_scale[0])
    # the same in two calls (whole 128-row blocks where they lie + the remainder), each into its rows / scale columns
    send2 = torch.zeros_like(send)
    qs2, ks2 = torch.zeros_like(qs), torch.zeros_like(ks)
    cut = 640
    (cq, ck), (eq, ek) = parallel._scale_counts(cut), parallel._scale_counts(n - cut)
    ops.quant_rows(stage[0], cut, None, False, world, send2[:, :cut, 0], qs2[:, :, :cq])
    ops.quant_rows(stage[0, cut:], n - cut, None, False, world, send2[:, cut:, 0], qs2[:, :, cq:cq + eq])
    ops.quant_rows(stage[1], cut, km, True, world, send2[:, :cut, 1], ks2[:, :, :ck])
    ops.quant_rows(stage[1, cut:], n - cut, km, True, world, send2[:, cut:, 1], ks2[:, :, ck:ck + ek])
    assert torch.equal(send2[:, :n], send[:, :n]) and torch.equal(qs2, qs) and torch.equal(ks2, ks)
    # the same through one base pointer per destination (here: separate local tensors; on several GPUs the peers'
    # symmetric-memory arenas), scales at a column offset of a wider per-head row, guard entries untouched
    dst = [torch.zeros((m, 2, hp, d), dtype=torch.int8, device="cuda") for _ in range(world)]
    wide_q = [torch.full((hp, nsq + 40), -7.0, device="cuda") for _ in range(world)]
    wide_k = [torch.full((hp, nsk + 12), -7.0, device="cuda") for _ in range(world)]
    ops.quant_rows_peers(stage[0], n, None, False, world, [t.data_ptr() for t in dst], d, 2 * hp * d,
                         [t.data_ptr() + 8 * 4 for t in wide_q], nsq + 40)
    ops.quant_rows_peers(stage[1], n, km, True, world, [t.data_ptr() + hp * d for t in dst], d, 2 * hp * d,
                         [t.data_ptr() + 4 * 4 for t in wide_k], nsk + 12)
    for p in range(world):
        assert torch.equal(dst[p][:n], send[p, :n])
        assert torch.equal(wide_q[p][:, 8:8 + nsq], qs[p]) and torch.equal(wide_k[p][:, 4:4 + nsk], ks[p])
        assert (wide_q[p][:, :8] == -7).all() and (wide_q[p][:, 8 + nsq:] == -7).all()
        assert (wide_k[p][:, :4] == -7).all() and (wide_k[p][:, 4 + nsk:] == -7).all()
    # V: conversion with a caller-provided range == the one-GPU per_channel_fp8 (codes, scales; padding zeroed)
    v8 = torch.full((world, hp, d, m), 0x7F, dtype=torch.uint8, device="cuda")
    v_scale = torch.empty((h, d), device="cuda")
    ops.v_convert(stage[2], n, st[1], world, v8, v_scale)
    n_stat = (n + 127) // 128 * 128
    r8, r_scale, _ = quant.per_channel_fp8(v, tensor_layout="NHD", scale_max=448.0, smooth_v=False, _n_stat=n_stat)
    npad = r8.size(-1)
    assert torch.equal(v8.reshape(h, d, m)[:, :, :npad], r8[0].view(torch.uint8).permute(1, 0, 2))
    assert torch.equal(v_scale, r_scale[0])
    v8b = torch.full_like(v8, 0x7F)
    ops.v_convert(stage[2], cut, st[1], world, v8b[..., :cut], v_scale)
    ops.v_convert(stage[2, cut:], n - cut, st[1], world, v8b[..., cut:], v_scale)
    assert torch.equal(v8b[..., :npad], v8[..., :npad])
    vdst = [torch.full((hp, d, 2 * m), 0x7F, dtype=torch.uint8, device="cuda") for _ in range(world)]
    v_scale2 = torch.empty_like(v_scale)
    ops.v_convert_peers(stage[2], n, st[1], world, [t.data_ptr() + m for t in vdst], d * 2 * m, 2 * m, v_scale2)
    for p in range(world):
        assert torch.equal(vdst[p][..., m:m + npad], v8[p][..., :npad]) and (vdst[p][..., :m] == 0x7F).all()
    assert torch.equal(v_scale2, v_scale)
    # the attention kernel on those buffers as the receiving rank sees them == the operator's own two-call path
    from sageattention_b200 import _lib, core
    o_ref = torch.empty((1, n, h, d), dtype=q.dtype, device="cuda")
    core._launch(True, q8, k8, r8, o_ref, q_scale, k_scale, r_scale, None, "NHD", False, _lib.GRAN_PER_THREAD, 32,
                 d ** -0.5, False)
    for p in range(world):
        o = ops.attention(send[p:p + 1, :n, 0], send[p:p + 1, :n, 1], v8[p], qs[p:p + 1], ks[p:p + 1],
                          v_scale[p * hp:(p + 1) * hp].contiguous(), q.dtype, False, d ** -0.5)
        assert torch.equal(o, o_ref[:, :, p * hp:(p + 1) * hp])
    # the output scattered by the kernel's epilogue (sage_attn_args.o_peer_*): rows [s * 400, (s + 1) * 400) of every head
    # group go to destination s's [400, H, D] buffer (here local tensors; tiles straddle the boundaries, the last
    # destination is only partly filled)
    rows = 400
    dests = [torch.full((rows, h, d), 7.0, dtype=q.dtype, device="cuda") for _ in range(3)]
    scratch = torch.empty(16 * 136, dtype=torch.uint8, device="cuda")     # SAGE_O_PEER_SCRATCH_BYTES
    for p in range(world):
        ops.attention(send[p:p + 1, :n, 0], send[p:p + 1, :n, 1], v8[p], qs[p:p + 1], ks[p:p + 1],
                      v_scale[p * hp:(p + 1) * hp].contiguous(), q.dtype, False, d ** -0.5,
                      o_peers=([t.data_ptr() + p * hp * d * 2 for t in dests], rows, d, h * d, scratch))
        torch.cuda.synchronize()
    got = torch.cat(dests)
    assert torch.equal(got[:n], o_ref[0])


def _q8_worker(rank, world, port, n, h, d, causal, q_out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import sageattention_b200 as sb
        from sageattention_b200 import parallel
        q, k, v = _q8_inputs(n, h, d)
        full = sb.sageattn(q, k, v, tensor_layout="NHD", is_causal=causal)
        nl = n // world
        sl = slice(rank * nl, (rank + 1) * nl)
        args = (q[:, sl].contiguous(), k[:, sl].contiguous(), v[:, sl].contiguous())
        o = parallel.ulysses_sageattn(*args, payload="int8", is_causal=causal)
        o2 = parallel.ulysses_sageattn(*args, payload="int8", is_causal=causal)     # buffers of a second call
        # the same with the quantisers storing into the peers' symmetric-memory arenas (no inbound collective): three
        # calls in a row (the arena is reused), the second on other data
        o3 = parallel.ulysses_sageattn(*args, payload="int8", transport="p2p", is_causal=causal)
        parallel.ulysses_sageattn(args[1], args[2], args[0], payload="int8", transport="p2p", is_causal=causal)
        o4 = parallel.ulysses_sageattn(*args, payload="int8", transport="p2p", is_causal=causal)
        torch.cuda.synchronize()
        o2 = o2 if (torch.equal(o3, o) and torch.equal(o4, o)) else o2 + 1   # any difference fails `repeat_same` below
        ref = torch.nn.functional.scaled_dot_product_attention(
            q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2),
            is_causal=causal).transpose(1, 2)[:, sl]
        f = lambda a, b_: torch.nn.functional.cosine_similarity(a.float().flatten(), b_.float().flatten(), dim=0).item()
        l1 = lambda a, b_: ((a.float() - b_.float()).abs().sum() / b_.float().abs().sum()).item()
        q_out.put((rank, tuple(o.shape), bool(torch.equal(o, o2)), f(o, full[:, sl]), l1(o, full[:, sl]),
                   l1(o, ref), l1(full[:, sl], ref), (o != full[:, sl]).float().mean().item()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,h,d,causal", [(2 * 1100, 8, 128, False), (2 * 1024, 4, 64, True), (2 * 1332, 4, 128, True)])
def test_ulysses_quantised_payload_two_gpus(n, h, d, causal):
    """payload="int8" on 2 GPUs against the one-GPU operator on the gathered inputs: Q / V codes and all scales are the
    one-GPU ones and K's mean can differ in the last bit of a channel (other summation order), so the outputs agree far
    inside the quantisation noise - and the error against fp32 attention is the one-GPU operator's."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = 29900 + os.getpid() % 1000 + d
    procs = [ctx.Process(target=_q8_worker, args=(r, 2, port, n, h, d, causal, q_out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, shape, repeat_same, cos, l1_full, l1_ref, l1_full_ref, frac_diff in res:
        print(f"rank {rank}: cos {cos:.7f} rel-L1 vs one GPU {l1_full:.2e}; vs fp32 {l1_ref:.4f} (one GPU {l1_full_ref:.4f}); "
              f"elements differing {frac_diff:.3%}")
        assert shape == (1, n // 2, h, d)
        assert repeat_same
        assert cos >= 0.9999 and l1_full <= 5e-3
        assert l1_ref <= 1.02 * l1_full_ref + 1e-4
