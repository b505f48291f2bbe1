"""Multi-GPU use of the operator.  The reference itself has no parallel code; its
example gets Ulysses sequence parallelism from xDiT by monkey-patching SDPA
(/root/reference/example/parallel_sageattn_cogvideo.py:8-52).  Every (batch, head) of the
operator is independent, so two layouts cover the named DiT configs:

* head sharding (`shard_heads`): rank r owns heads [r*H/P, (r+1)*H/P) - no communication at all;
* Ulysses (`ulysses_sageattn`): activations arrive sequence-sharded [B, N/P, H, D]; ONE all-to-all
  scatters heads / gathers the sequence for q, k, v together, the operator runs on H/P heads of the
  FULL sequence (so K smoothing, scales and therefore numerics are identical to one GPU), and one
  all-to-all brings O back to [B, N/P, H, D].  With B = 1 the rank's heads can be pipelined in groups
  (`head_chunks`; measured neutral with NCCL, whose kernels need the SMs the persistent attention grid
  holds - DESIGN 5).  NCCL over NVLink/NVSwitch on GPUs; the same code runs
  over gloo on CPU tensors for the host-logic tests (with a stand-in attention function).
"""
from typing import Callable, Optional

import torch
import torch.distributed as dist


def shard_heads(x: torch.Tensor, rank: int, world: int, tensor_layout: str = "HND") -> torch.Tensor:
    """View of this rank's heads of a [B,H,N,D] ("HND") or [B,N,H,D] ("NHD") tensor."""
    dim = 1 if tensor_layout == "HND" else 2
    h = x.size(dim)
    if h % world:
        raise ValueError(f"num_heads {h} is not divisible by world size {world}")
    return x.narrow(dim, rank * (h // world), h // world)


def _seq_to_head_shard(x: torch.Tensor, group, world: int) -> torch.Tensor:
    """[T, B, n_loc, H, D] (sequence-sharded) -> [T, B, n_loc*P, H/P, D] (head-sharded)."""
    t, b, n, h, d = x.shape
    hp = h // world
    # chunk p of the send buffer = heads [p*hp, (p+1)*hp) -> goes to rank p
    send = x.reshape(t, b, n, world, hp, d).permute(3, 0, 1, 2, 4, 5).contiguous()
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    # chunk p of the receive buffer = sequence block p (from rank p) of my heads
    return recv.permute(1, 2, 0, 3, 4, 5).reshape(t, b, world * n, hp, d)


def _head_to_seq_shard(x: torch.Tensor, group, world: int) -> torch.Tensor:
    """[B, n_loc*P, H/P, D] (head-sharded) -> [B, n_loc, H, D] (sequence-sharded)."""
    b, nt, hp, d = x.shape
    n = nt // world
    send = x.reshape(b, world, n, hp, d).permute(1, 0, 2, 3, 4).contiguous()
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    # chunk p = heads of rank p for my sequence block
    return recv.permute(1, 2, 0, 3, 4).reshape(b, n, world * hp, d)


def _a2a_heads_in(x: torch.Tensor, lo: int, hi: int, group, world: int) -> torch.Tensor:
    """Batch-1 path in: heads [lo, hi) of EVERY rank's head block, [1, n_loc, H, D] ->
    [1, n_loc*P, hi-lo, D].  One copy (the head gather into the send buffer); the receive buffer,
    chunked by source rank = by sequence block, already IS the head-sharded tensor."""
    _, n, h, d = x.shape
    hp = h // world
    send = torch.empty((world, n, hi - lo, d), dtype=x.dtype, device=x.device)
    send.copy_(x.view(n, world, hp, d)[:, :, lo:hi].transpose(0, 1))
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    return recv.view(1, world * n, hi - lo, d)


def _a2a_heads_in_qkv(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, lo: int, hi: int, group, world: int):
    """q, k, v of equal shape (no GQA): ONE all-to-all for all three instead of three.  The send buffer is laid out
    [rank, token, (q|k|v), head, d], so in the receive buffer every tensor is a strided [1, n_loc*P, hi-lo, D] view with
    a uniform token stride of 3*(hi-lo)*D - which the kernels take as it is (their TMA maps are built from the real
    strides).  Same number of gather copies as the three-call path (one per tensor), two collective launches fewer."""
    _, n, h, d = q.shape
    hp = h // world
    hc = hi - lo
    send = torch.empty((world, n, 3, hc, d), dtype=q.dtype, device=q.device)
    for i, x in enumerate((q, k, v)):
        send[:, :, i].copy_(x.view(n, world, hp, d)[:, :, lo:hi].transpose(0, 1))
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    r = recv.view(1, world * n, 3, hc, d)
    return r[:, :, 0], r[:, :, 1], r[:, :, 2]


def _a2a_heads_out(o: torch.Tensor, group, world: int) -> torch.Tensor:
    """Batch-1 path back: [1, n_loc*P, hc, D] (contiguous: it is its own send buffer, chunk p = sequence
    block p) -> [P, n_loc, hc, D], chunk p = these heads as computed by rank p."""
    _, nt, hc, d = o.shape
    recv = torch.empty((world, nt // world, hc, d), dtype=o.dtype, device=o.device)
    dist.all_to_all_single(recv, o.view(world, nt // world, hc, d), group=group)
    return recv


_COMM_STREAMS = {}
# B = 1, equal q / k / v shapes: one inbound all-to-all for q, k and v (strided views of one receive buffer) instead of three
FUSED_QKV_A2A = True


def _ulysses_b1(q, k, v, group, world, attn_fn, head_chunks, kwargs):
    """B = 1 (the DiT case).  The rank's heads are processed in `head_chunks` groups: all inbound
    all-to-alls are queued on a side stream, attention on group c starts as soon as ITS q/k/v have
    arrived and runs under the transfers of group c+1, and its output travels back under the attention of
    the next group.  4 tensor copies in total (11 in the generic path)."""
    n, h, d = q.shape[1:]
    hk = k.size(2)
    hp, hkp = h // world, hk // world
    g = h // hk
    chunks = max(1, min(head_chunks, hkp))
    bounds = [(hkp * c) // chunks for c in range(chunks + 1)]  # kv-head ranges; q heads follow (GQA)
    cuda = q.is_cuda
    if cuda:
        cur = torch.cuda.current_stream(q.device)
        comm = _COMM_STREAMS.get(q.device)
        if comm is None:
            comm = _COMM_STREAMS[q.device] = torch.cuda.Stream(q.device)
        comm.wait_stream(cur)
    ins = []
    for c in range(chunks):
        lo, hi = bounds[c], bounds[c + 1]
        fused = FUSED_QKV_A2A and g == 1 and k.shape == q.shape and v.shape == q.shape
        if cuda:
            with torch.cuda.stream(comm):
                if fused:
                    t = _a2a_heads_in_qkv(q, k, v, lo, hi, group, world)
                else:
                    t = tuple(_a2a_heads_in(x, a, b_, group, world)
                              for x, a, b_ in ((q, lo * g, hi * g), (k, lo, hi), (v, lo, hi)))
                ev = torch.cuda.Event()
                ev.record(comm)
            for x in t:
                x.record_stream(cur)
        elif fused:
            t = _a2a_heads_in_qkv(q, k, v, lo, hi, group, world)
            ev = None
        else:
            t = (_a2a_heads_in(q, lo * g, hi * g, group, world), _a2a_heads_in(k, lo, hi, group, world),
                 _a2a_heads_in(v, lo, hi, group, world))
            ev = None
        ins.append((t, ev))
    out = torch.empty((1, n, h, d), dtype=q.dtype, device=q.device)
    outs = []
    for c, ((qh, kh, vh), ev) in enumerate(ins):
        if cuda:
            cur.wait_event(ev)
        o = attn_fn(qh, kh, vh, tensor_layout="NHD", **kwargs).contiguous()
        if cuda:
            done = torch.cuda.Event()
            done.record(cur)
            o.record_stream(comm)
            with torch.cuda.stream(comm):
                comm.wait_event(done)
                r = _a2a_heads_out(o, group, world)
            r.record_stream(cur)
        else:
            r = _a2a_heads_out(o, group, world)
        outs.append(r)
    if cuda:
        cur.wait_stream(comm)
    ov = out.view(n, world, hp, d)
    for c, r in enumerate(outs):
        ov[:, :, bounds[c] * g:bounds[c + 1] * g].copy_(r.transpose(0, 1))  # head interleave of the received chunks
    return out


# ---------------------------------------------------------------------------------------------------
# Batch-1 Ulysses over NVLink peer memory (no NCCL kernel on the data path).  NCCL's all-to-all runs on SMs,
# and the persistent attention grid holds all of them, so NCCL transfers cannot overlap the attention of
# another head group.  Here every rank owns a symmetric-memory arena (torch.distributed._symmetric_memory:
# CUDA IPC / fabric handles exchanged once); a rank PUSHES its head-group chunks straight into the peers'
# arenas with contiguous device-to-device copies, which the copy engines execute - no SM needed - and a
# signal-pad barrier tells the receivers that everything has landed.
_ARENAS = {}


def _arena(nbytes: int, device: torch.device, group):
    """(handle, capacity) of this rank's symmetric arena of at least `nbytes` (same on every rank: the size
    is a function of the call's shapes, so all ranks re-allocate together)."""
    import torch.distributed._symmetric_memory as symm
    key = (device.index, group.group_name)
    cur = _ARENAS.get(key)
    if cur is None or cur[1] < nbytes:
        cap = max(nbytes, 1 << 20)
        buf = symm.empty(cap, dtype=torch.uint8, device=device)
        cur = (symm.rendezvous(buf, group), cap, buf)
        _ARENAS[key] = cur
    return cur[0], cur[1]


def _ulysses_b1_p2p(q, k, v, group, world, attn_fn, head_chunks, kwargs):
    rank = dist.get_rank(group)
    n, h, d = q.shape[1:]
    hk = k.size(2)
    hp, hkp = h // world, hk // world
    g = h // hk
    es = q.element_size()
    chunks = max(1, min(head_chunks, hkp))
    bounds = [(hkp * c) // chunks for c in range(chunks + 1)]
    # arena layout (bytes), identical on every rank: per chunk [q | k | v | o_back], each [P (source rank), n, heads, d]
    offs, total = [], 0
    for c in range(chunks):
        hc_k = bounds[c + 1] - bounds[c]
        sizes = [world * n * hc_k * g * d * es, world * n * hc_k * d * es, world * n * hc_k * d * es,
                 world * n * hc_k * g * d * es]
        o4 = []
        for sz in sizes:
            o4.append(total)
            total += (sz + 255) // 256 * 256
        offs.append(o4)
    hdl, _ = _arena(total, q.device, group)
    dt = q.dtype

    def peer(p, byte_off, heads):  # rank p's slot for the rows this rank sends: [n, heads, d]
        return hdl.get_buffer(p, (n, heads, d), dt, (byte_off + rank * n * heads * d * es) // es)

    def mine(byte_off, heads):     # this rank's whole buffer: [P, n, heads, d]
        return hdl.get_buffer(rank, (world, n, heads, d), dt, byte_off // es)

    cur = torch.cuda.current_stream(q.device)
    comm = _COMM_STREAMS.get(q.device)
    if comm is None:
        comm = _COMM_STREAMS[q.device] = torch.cuda.Stream(q.device)
    comm.wait_stream(cur)
    arrived = []
    with torch.cuda.stream(comm):
        hdl.barrier()  # every rank has finished reading its arena in the previous call
        for c in range(chunks):
            lo, hi = bounds[c], bounds[c + 1]
            for x, a, b_, off in ((q, lo * g, hi * g, offs[c][0]), (k, lo, hi, offs[c][1]), (v, lo, hi, offs[c][2])):
                hpx = x.size(2) // world
                send = torch.empty((world, n, b_ - a, d), dtype=dt, device=x.device)
                send.copy_(x.view(n, world, hpx, d)[:, :, a:b_].transpose(0, 1))   # the one gather copy (local)
                for p in range(world):
                    peer(p, off, b_ - a).copy_(send[p])                           # contiguous: copy engine, NVLink
            hdl.barrier()  # all pushes of this head group have landed everywhere
            ev = torch.cuda.Event()
            ev.record(comm)
            arrived.append(ev)
    out = torch.empty((1, n, h, d), dtype=dt, device=q.device)
    ov = out.view(n, world, hp, d)
    back = []
    for c in range(chunks):
        lo, hi = bounds[c], bounds[c + 1]
        cur.wait_event(arrived[c])
        qh = mine(offs[c][0], (hi - lo) * g).view(1, world * n, (hi - lo) * g, d)
        kh = mine(offs[c][1], hi - lo).view(1, world * n, hi - lo, d)
        vh = mine(offs[c][2], hi - lo).view(1, world * n, hi - lo, d)
        o = attn_fn(qh, kh, vh, tensor_layout="NHD", **kwargs).contiguous()
        done = torch.cuda.Event()
        done.record(cur)
        o.record_stream(comm)
        with torch.cuda.stream(comm):
            comm.wait_event(done)
            ob = o.view(world, n, (hi - lo) * g, d)   # chunk p = sequence block p: goes to rank p
            for p in range(world):
                peer(p, offs[c][3], (hi - lo) * g).copy_(ob[p])
            hdl.barrier()
            ev = torch.cuda.Event()
            ev.record(comm)
            back.append(ev)
    for c in range(chunks):
        lo, hi = bounds[c], bounds[c + 1]
        cur.wait_event(back[c])
        ov[:, :, lo * g:hi * g].copy_(mine(offs[c][3], (hi - lo) * g).transpose(0, 1))
    comm.wait_stream(cur)  # the arena is read by `cur` until here; the next call's first barrier follows on comm
    return out


def ulysses_sageattn(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, group=None,
                     attn_fn: Optional[Callable] = None, head_chunks: int = 1, transport: str = "nccl",
                     comm_sms: int = 0, **kwargs) -> torch.Tensor:
    """Sequence-parallel attention.  q, k, v: this rank's sequence block, "NHD" layout
    [B, N/P, H, D] (all ranks hold equally long blocks; pad the sequence to a multiple of P first).
    Returns this rank's block of the output, [B, N/P, H, D].  kwargs go to the attention function
    (default: sageattn); `tensor_layout` is fixed to "NHD" and `return_lse` is not supported here.
    `transport`: "nccl" (all_to_all_single) or, with B = 1 on GPUs of one node, "p2p": chunks are pushed into
    the peers' symmetric-memory arenas by the copy engines, which (unlike NCCL's kernels) run under the
    persistent attention grid of another head group.
    `head_chunks`: with B = 1, how many groups the rank's heads are split into so that the all-to-alls of
    one group overlap the attention of another (1 = no overlap).
    `comm_sms`: with head_chunks > 1 and NCCL, how many SMs the persistent attention grid leaves free, so that
    NCCL's all-to-all kernels (which need SMs) really run next to the attention of another group."""
    if kwargs.pop("return_lse", False):
        raise ValueError("ulysses_sageattn does not return the LSE")
    kwargs.pop("tensor_layout", None)
    if attn_fn is None:
        from .core import sageattn as attn_fn
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return attn_fn(q, k, v, tensor_layout="NHD", **kwargs)
    h = q.size(2)
    if h % world or k.size(2) % world:
        raise ValueError(f"head counts ({h}, {k.size(2)}) must be divisible by the world size {world}")
    if q.size(0) == 1 and k.size(0) == 1 and all(t.is_contiguous() for t in (q, k, v)):
        # the DiT case (B = 1): communication pipelined against attention over groups of heads
        if transport == "p2p" and (k.size(1) != q.size(1) or v.size(1) != k.size(1)):
            # the peer-memory path lays its arena out for ONE sequence-shard length (q, k, v alike); cross attention
            # with another kv length takes the NCCL path, which derives the length per tensor (round-1 ADVICE)
            transport = "nccl"
        if transport == "p2p":
            if not q.is_cuda:
                raise ValueError("transport='p2p' needs CUDA tensors")
            return _ulysses_b1_p2p(q, k, v, group if group is not None else dist.group.WORLD, world, attn_fn,
                                   head_chunks, kwargs)
        if comm_sms > 0 and head_chunks > 1 and q.is_cuda:
            from . import core
            sms = torch.cuda.get_device_properties(q.device).multi_processor_count
            prev, core.GRID_LIMIT = core.GRID_LIMIT, max(1, sms - comm_sms)
            try:
                return _ulysses_b1(q, k, v, group, world, attn_fn, head_chunks, kwargs)
            finally:
                core.GRID_LIMIT = prev
        return _ulysses_b1(q, k, v, group, world, attn_fn, head_chunks, kwargs)
    if k.size(2) == h and k.shape == q.shape:
        qkv = _seq_to_head_shard(torch.stack((q, k, v)), group, world)  # one collective for all three
        qh, kh, vh = qkv[0], qkv[1], qkv[2]
    else:  # GQA / cross attention: q and k/v differ in shape
        qh = _seq_to_head_shard(q.unsqueeze(0), group, world)[0]
        kv = _seq_to_head_shard(torch.stack((k, v)), group, world)
        kh, vh = kv[0], kv[1]
    o = attn_fn(qh, kh, vh, tensor_layout="NHD", **kwargs)
    return _head_to_seq_shard(o.contiguous(), group, world)
