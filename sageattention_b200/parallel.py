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
* Ulysses with a quantised payload (`payload="int8"`): the pre-pass runs on the sequence shards BEFORE the
  exchange, INT8 / e4m3 codes and scales travel instead of 16-bit activations, and the receiving rank goes straight into
  the attention kernel (`_ulysses_b1_q8`, NCCL).  With `transport="p2p"` (`_ulysses_b1_q8_p2p`) there is no collective on
  the data path at all: the quantiser launches store into the destination ranks' symmetric-memory arenas over NVLink and
  the attention kernel's TMA-store epilogue scatters the output rows to the ranks that own them (Wan2.1 layer on 8 GPUs:
  10.5 ms, 0.97 of linear, against 11.2 ms / 0.91 for the 16-bit NCCL path).
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
_ARENAS_RETIRED = []


def _arena(nbytes: int, device: torch.device, group):
    """(handle, capacity) of this rank's symmetric arena of at least `nbytes` (same on every rank: the size
    is a function of the call's shapes, so all ranks re-allocate together)."""
    import torch.distributed._symmetric_memory as symm
    key = (device.index, group.group_name)
    cur = _ARENAS.get(key)
    if cur is None or cur[1] < nbytes:
        if cur is not None:
            # kernels of the previous call (here or on a peer) may still be reading / writing the old arena: it is kept
            # alive instead of being unmapped under them (growth is rare: the size is a function of the call's shapes)
            _ARENAS_RETIRED.append(cur)
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


# ---------------------------------------------------------------------------------------------------
# Quantised payload: the pre-pass runs on the sequence shards BEFORE the all-to-all, so INT8 Q / K codes,
# e4m3 V codes and their scales travel (1 byte per element instead of 2) and the rank that receives them goes straight
# into the attention kernel.  What makes the codes those of the one-GPU call:
#   * quantisation groups are blocks of the GLOBAL sequence (128 q rows, 64 k rows, 16-token V permutation groups), and
#     sequence shards are not aligned to them (Wan2.1: 75 600 / 8 = 9 450 rows).  So ownership is re-cut on 128-row
#     boundaries: rank r quantises global rows [r*M, (r+1)*M), M = ceil128(n_loc); the rows it lacks - at most
#     (r+1)*(M - n_loc) < 128*P - come from rank r+1 in one small neighbour exchange (the "halo");
#   * K's mean and V's channel range are global: every rank reduces its own rows (sage_col_sum_absmax), the [2, H, D]
#     partials are all-gathered and folded in rank order on every rank (deterministic; Q and V codes and all scales are
#     bit-identical to one GPU, K's mean can differ from the one-GPU summation order in the last bit of a channel);
#   * the quantisers write straight into the send layout: with B = 1 the batch dimension of the C-ABI tensors is free,
#     so "batch" = destination rank (stride = its head group) and one launch per tensor produces [P, M, heads/P, D].
# Inbound collectives: one for the Q / K codes (token-major, read in place by the attention kernel through strided TMA
# maps), one for V^T codes + the Q / K scales.  The way back is the 16-bit output as before.
class _Q8CudaOps:
    """The device half of the quantised-payload path, over the C ABI (tests/test_parallel_cpu.py substitutes a CPU
    stand-in with the same methods to exercise the host logic over gloo)."""
    scale_max = 448.0   # sageattn: pv_accum_dtype "fp32+fp32" (/root/reference/sageattention/core.py:818-820)

    def stats(self, k, v):
        """[2, H, D] f32: column sums of this shard's K rows, max|v| per channel of its V rows."""
        from . import _lib
        from .quant import _dtype_code, _on_device, _stream
        lib = _lib.load()
        _, n, h, d = k.shape
        out = torch.empty((2, h, d), dtype=torch.float32, device=k.device)
        ws = torch.empty(lib.sage_seq_mean_workspace_bytes(1, h, n, d), dtype=torch.uint8, device=k.device)
        with _on_device(k.device):
            for x, so, ao in ((k, out[0].data_ptr(), None), (v, None, out[1].data_ptr())):
                x4 = _lib.Tensor4(x.data_ptr(), x.stride(0), x.stride(2), x.stride(1))
                _lib.check(lib.sage_col_sum_absmax(x4, so, ao, 1, h, n, d, _dtype_code(x), ws.data_ptr(), _stream(x)))
        return out

    def mean_from_sum(self, ksum, n_total, dtype):
        # sage_seq_mean's arithmetic (torch's mean): f32 sum * (float(outputs) / float(elements)), rounded to the dtype
        h, d = ksum.shape
        factor = torch.tensor(float(h * d), dtype=torch.float32) / torch.tensor(float(h * d) * n_total, dtype=torch.float32)
        return (ksum * factor.item()).to(dtype)

    def quant_rows(self, x, rows, km, is_key, world, out, scale):
        """x: [>= rows, H, D] rows starting on a 128-row boundary of the global sequence.  out: int8 [P, >= rows, H/P, D]
        view (any strides, rows contiguous), scale: f32 [P, H/P, entries] view with any head stride - per_thread codes
        and scales of heads [p*H/P, (p+1)*H/P) go to out[p] / scale[p]."""
        from . import _lib
        from .quant import _dtype_code, _on_device, _stream
        lib = _lib.load()
        m, h, d = x.shape
        hp = h // world
        assert scale.stride(2) == 1 and scale.stride(0) == hp * scale.stride(1)
        x4 = _lib.Tensor4(x.data_ptr(), hp * x.stride(1), x.stride(1), x.stride(0))
        o4 = _lib.Tensor4(out.data_ptr(), out.stride(0), out.stride(2), out.stride(1))
        with _on_device(x.device):
            _lib.check(lib.sage_quant_per_thread_int8_strided(
                x4, km.data_ptr() if km is not None else None, o4, scale.data_ptr(), world, hp, rows, d, _dtype_code(x),
                1 if is_key else 0, 64 if is_key else 32, scale.stride(1), _stream(x)))

    def v_convert(self, x, rows, absmax, world, out, v_scale):
        """x: [>= rows, H, D] rows of V starting on a 128-row boundary.  out: uint8 [P, H/P, D, >= ceil128(rows)] view;
        v_scale: f32 [H, D] (absmax / scale_max)."""
        from . import _lib
        from .quant import _dtype_code, _on_device, _stream
        lib = _lib.load()
        m, h, d = x.shape
        hp = h // world
        x4 = _lib.Tensor4(x.data_ptr(), hp * x.stride(1), x.stride(1), x.stride(0))
        n_stat = (rows + 127) // 128 * 128
        with _on_device(x.device):
            _lib.check(lib.sage_v_fp8_from_absmax(x4, out.data_ptr(), out.stride(0), out.stride(1), out.stride(2),
                                                  absmax.data_ptr(), v_scale.data_ptr(), world, hp, rows, n_stat, d,
                                                  _dtype_code(x), self.scale_max, _stream(x)))

    def quant_rows_peers(self, x, rows, km, is_key, world, out_ptrs, out_stride_h, out_stride_n, scale_ptrs,
                         scale_per_bh):
        """quant_rows with one output / scale base address per destination rank (ints: addresses inside the peers'
        symmetric-memory arenas): the stores are the all-to-all."""
        import ctypes
        from . import _lib
        from .quant import _dtype_code, _on_device, _stream
        lib = _lib.load()
        m, h, d = x.shape
        hp = h // world
        x4 = _lib.Tensor4(x.data_ptr(), hp * x.stride(1), x.stride(1), x.stride(0))
        op = (ctypes.c_void_p * world)(*out_ptrs)
        sp = (ctypes.c_void_p * world)(*scale_ptrs)
        with _on_device(x.device):
            _lib.check(lib.sage_quant_per_thread_int8_peers(
                x4, km.data_ptr() if km is not None else None, op, out_stride_h, out_stride_n, sp, world, hp, rows, d,
                _dtype_code(x), 1 if is_key else 0, 64 if is_key else 32, scale_per_bh, _stream(x)))

    def v_convert_peers(self, x, rows, absmax, world, out_ptrs, out_stride_h, out_stride_d, v_scale):
        import ctypes
        from . import _lib
        from .quant import _dtype_code, _on_device, _stream
        lib = _lib.load()
        m, h, d = x.shape
        hp = h // world
        x4 = _lib.Tensor4(x.data_ptr(), hp * x.stride(1), x.stride(1), x.stride(0))
        op = (ctypes.c_void_p * world)(*out_ptrs)
        n_stat = (rows + 127) // 128 * 128
        with _on_device(x.device):
            _lib.check(lib.sage_v_fp8_from_absmax_peers(x4, op, out_stride_h, out_stride_d, absmax.data_ptr(),
                                                        v_scale.data_ptr(), world, hp, rows, n_stat, d, _dtype_code(x),
                                                        self.scale_max, _stream(x)))

    def attention(self, q8, k8, v8, q_scale, k_scale, v_scale, out_dtype, is_causal, sm_scale, o_peers=None):
        """q8 / k8: int8 [1, N, h, D] (strided), v8: uint8 [h, D, Npad]; -> o [1, N, h, D].  With `o_peers` (see
        core._launch) the rows go to the destination ranks' buffers instead and `o` is a one-tile dummy."""
        from . import _lib, core
        _, n, h, d = q8.shape
        # (with o_peers nothing is stored through `o`: its tensor map only has to be constructible)
        o = torch.empty((1, n if o_peers is None else 128, h, d), dtype=out_dtype, device=q8.device)
        v4 = v8.view(torch.float8_e4m3fn).unsqueeze(0).transpose(1, 2)   # "NHD" form of v_fp8: [B, D, H, Npad]
        core._launch(True, q8, k8, v4, o, q_scale, k_scale, v_scale, None, "NHD", is_causal, _lib.GRAN_PER_THREAD, 32,
                     sm_scale, False, o_peers=o_peers)
        return o


def q8_geometry(n: int, world: int):
    """Ownership of the quantised-payload path for shards of n rows: (M, [rows owned by rank 0..P-1]) or None when the
    shards are too short for a one-neighbour halo."""
    m = (n + 127) // 128 * 128
    if world * (m - n) > n:
        return None
    total = world * n
    own = [max(0, min((r + 1) * m, total) - r * m) for r in range(world)]
    return m, own


def _scale_counts(rows: int):
    """per_thread scale entries per head for `rows` rows: (query: 8 per 32-row warp tile of whole 128-row blocks,
    key: 4 per 64-row block) - include/sage_b200.h, sage_quant_per_thread_int8."""
    return (rows + 127) // 128 * 32, (rows + 63) // 64 * 4


def _ulysses_b1_q8(q, k, v, group, world, ops, is_causal=False, sm_scale=None):
    rank = dist.get_rank(group)
    _, n, h, d = q.shape
    hp = h // world
    m, own = q8_geometry(n, world)
    total = world * n
    mine = own[rank]
    dev, dt = q.device, q.dtype
    peer = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)

    # 1. the rows this rank owns = the tail of its shard (from local row `give`) + the head of the next rank's (the
    #    halo).  Whole 128-row blocks that lie inside the shard are quantised where they are; the rest (< 128 local rows
    #    + the halo) goes through a small staging buffer.  The neighbour exchange is issued first: it runs under the
    #    statistics kernels.
    give = rank * (m - n)                 # my first rows belong to rank - 1
    keep = n - give                       # local rows I own
    halo = mine - keep                    # rows of rank + 1 I own
    main = keep // 128 * 128 if halo > 0 else keep   # (no halo: the shard's tail is the sequence's, one direct call)
    edge = mine - main                    # rows through the staging buffer
    stage = torch.empty((3, max(edge, 1), h, d), dtype=dt, device=dev)
    ops_p2p = []
    for i, x in enumerate((q, k, v)):
        if give > 0:
            ops_p2p.append(dist.P2POp(dist.isend, x[0, :give], peer(rank - 1), group))
        if halo > 0:
            ops_p2p.append(dist.P2POp(dist.irecv, stage[i, keep - main:edge], peer(rank + 1), group))
    works = dist.batch_isend_irecv(ops_p2p) if ops_p2p else []

    # 2. partial statistics of the local rows -> global K mean and V range (same fold on every rank)
    part = ops.stats(k, v)
    allp = torch.empty((world,) + tuple(part.shape), dtype=part.dtype, device=dev)
    gathered = dist.all_gather(list(allp.unbind(0)), part, group=group, async_op=True)

    # 3. quantise into the send layout (Q first: it needs no statistics and runs under the all-gather)
    nsq, nsk = _scale_counts(m)           # per-head scale capacity of a chunk (M rows)
    cv = hp * d * m + hp * (nsq + nsk) * 4          # bytes per destination of the V / scales buffer
    qk_send = torch.empty((world, m, 2, hp, d), dtype=torch.int8, device=dev)
    vs_send = torch.empty((world, cv), dtype=torch.uint8, device=dev)
    v_send = vs_send[:, :hp * d * m].view(world, hp, d, m)
    tail = vs_send[:, hp * d * m:].view(torch.float32)          # [P, hp * nsq | hp * nsk]
    # (the quantisers address scales as [batch = destination, head, entry] with a dense batch stride: own arrays, copied
    # behind the V codes of every destination afterwards)
    qs_send = torch.empty((world, hp, nsq), dtype=torch.float32, device=dev)
    ks_send = torch.empty((world, hp, nsk), dtype=torch.float32, device=dev)
    v_scale = torch.empty((h, d), dtype=torch.float32, device=dev)
    mq, mk = _scale_counts(main)
    eq, ek = _scale_counts(edge)
    if main > 0:
        ops.quant_rows(q[0, give:give + main], main, None, False, world, qk_send[:, :main, 0], qs_send[:, :, :mq])
    gathered.wait()
    km = ops.mean_from_sum(allp[:, 0].sum(dim=0), total, dt)
    vmax = allp[:, 1].amax(dim=0)
    if main > 0:
        ops.quant_rows(k[0, give:give + main], main, km, True, world, qk_send[:, :main, 1], ks_send[:, :, :mk])
    if edge > 0:
        for i, x in enumerate((q, k, v)):
            if keep > main:
                stage[i, :keep - main].copy_(x[0, give + main:])
        for w in works:
            w.wait()
        ops.quant_rows(stage[0], edge, None, False, world, qk_send[:, main:, 0], qs_send[:, :, mq:mq + eq])
        ops.quant_rows(stage[1], edge, km, True, world, qk_send[:, main:, 1], ks_send[:, :, mk:mk + ek])
    else:
        for w in works:
            w.wait()

    # 4. the two inbound collectives; the codes of Q and K leave while V is still being converted
    qk_recv = torch.empty_like(qk_send)
    vs_recv = torch.empty_like(vs_send)
    qk_sent = dist.all_to_all_single(qk_recv, qk_send, group=group, async_op=True)
    if main > 0:
        ops.v_convert(v[0, give:give + main], main, vmax, world, v_send[..., :(main + 127) // 128 * 128], v_scale)
    if edge > 0:
        ops.v_convert(stage[2], edge, vmax, world, v_send[..., main:], v_scale)
    tail[:, :hp * nsq].copy_(qs_send.view(world, hp * nsq))
    tail[:, hp * nsq:].copy_(ks_send.view(world, hp * nsk))
    dist.all_to_all_single(vs_recv, vs_send, group=group)
    qk_sent.wait()

    # 5. this rank's heads of the whole sequence: codes in place, V^T and the scales re-assembled by source rank (every
    #    source but the last non-empty one owns M rows)
    qk = qk_recv.view(1, world * m, 2, hp, d)
    q8, k8 = qk[:, :total, 0], qk[:, :total, 1]
    v8 = torch.empty((hp, d, world * m), dtype=torch.uint8, device=dev)
    v8.view(hp, d, world, m).copy_(vs_recv[:, :hp * d * m].view(world, hp, d, m).permute(1, 2, 0, 3))
    tq, tk = _scale_counts(total)
    rtail = vs_recv[:, hp * d * m:].view(torch.float32)
    rq = rtail[:, :hp * nsq].view(world, hp, nsq)
    rk = rtail[:, hp * nsq:].view(world, hp, nsk)
    nfull = sum(1 for o_ in own if o_ == m)
    q_scale = torch.empty((1, hp, tq), dtype=torch.float32, device=dev)
    k_scale = torch.empty((1, hp, tk), dtype=torch.float32, device=dev)
    q_scale[0, :, :nfull * nsq].view(hp, nfull, nsq).copy_(rq[:nfull].permute(1, 0, 2))
    k_scale[0, :, :nfull * nsk].view(hp, nfull, nsk).copy_(rk[:nfull].permute(1, 0, 2))
    if nfull < world and own[nfull] > 0:
        cq, ck = _scale_counts(own[nfull])
        q_scale[0, :, nfull * nsq:].copy_(rq[nfull, :, :cq])
        k_scale[0, :, nfull * nsk:].copy_(rk[nfull, :, :ck])
        assert nfull * nsq + cq == tq and nfull * nsk + ck == tk, "scale bookkeeping"
    if sm_scale is None:
        sm_scale = d ** -0.5
    o = ops.attention(q8, k8, v8, q_scale, k_scale, v_scale[rank * hp:(rank + 1) * hp].contiguous(), dt, is_causal,
                      sm_scale)

    # 6. the way back: 16-bit output, one all-to-all + the head interleave
    r_ = _a2a_heads_out(o, group, world)
    out = torch.empty((1, n, h, d), dtype=dt, device=dev)
    out.view(n, world, hp, d).copy_(r_.transpose(0, 1))
    return out


def _ulysses_b1_q8_p2p(q, k, v, group, world, ops, is_causal=False, sm_scale=None, peer_out=True):
    """_ulysses_b1_q8 with NO inbound collective: every rank owns a symmetric-memory arena laid out as the attention
    kernel wants its inputs ([P, M, 2, H/P, D] Q / K codes by source rank, V^T [H/P, D, P*M], scales [H/P, entries]), and
    the quantisers of every rank store their codes and scales straight into the destinations' arenas over NVLink
    (sage_quant_per_thread_int8_peers / sage_v_fp8_from_absmax_peers: one base pointer per destination) - the launch is
    quantiser and all-to-all in one.  Two signal-pad barriers frame the stores (everybody has finished reading the
    previous call's arena / everything has landed).  `peer_out`: the way back goes through the attention kernel's
    epilogue as well (sage_attn_args.o_peer_*: rows of destination s are TMA-stored into s's arena), so the only NCCL
    traffic left is the small statistics all-gather and the halo; False = 16-bit output through NCCL as in _ulysses_b1_q8."""
    group = group if group is not None else dist.group.WORLD
    rank = dist.get_rank(group)
    _, n, h, d = q.shape
    hp = h // world
    m, own = q8_geometry(n, world)
    total = world * n
    mine = own[rank]
    dev, dt = q.device, q.dtype
    peer = lambda r: dist.get_global_rank(group, r)
    nsq, nsk = _scale_counts(m)
    tq, tk = _scale_counts(total)
    al = lambda x: (x + 255) // 256 * 256
    off_v = al(world * m * 2 * hp * d)
    off_qs = off_v + al(hp * d * world * m)
    off_ks = off_qs + al(hp * tq * 4)
    off_o = off_ks + al(hp * tk * 4)               # this rank's output rows, [n, H, D], written by every rank's epilogue
    es = q.element_size()
    nbytes = off_o + al(n * h * d * es) if peer_out else off_o
    hdl, _ = _arena(nbytes, dev, group)
    base = [hdl.get_buffer(p, (nbytes,), torch.uint8, 0).data_ptr() for p in range(world)]

    give = rank * (m - n)
    keep = n - give
    halo = mine - keep
    main = keep // 128 * 128 if halo > 0 else keep
    edge = mine - main
    stage = torch.empty((3, max(edge, 1), h, d), dtype=dt, device=dev)
    ops_p2p = []
    for i, x in enumerate((q, k, v)):
        if give > 0:
            ops_p2p.append(dist.P2POp(dist.isend, x[0, :give], peer(rank - 1), group))
        if halo > 0:
            ops_p2p.append(dist.P2POp(dist.irecv, stage[i, keep - main:edge], peer(rank + 1), group))
    works = dist.batch_isend_irecv(ops_p2p) if ops_p2p else []
    part = ops.stats(k, v)
    allp = torch.empty((world,) + tuple(part.shape), dtype=part.dtype, device=dev)
    gathered = dist.all_gather(list(allp.unbind(0)), part, group=group, async_op=True)

    hdl.barrier()      # every rank's attention of the previous call has finished reading its arena
    row = 2 * hp * d   # bytes between consecutive tokens of the Q / K code buffer
    v_scale = torch.empty((h, d), dtype=torch.float32, device=dev)
    mq, mk = _scale_counts(main)

    def code_ptrs(is_key, row0):
        return [b_ + rank * m * row + row0 * row + (hp * d if is_key else 0) for b_ in base]

    def scale_ptrs(is_key, col0):
        return [b_ + (off_ks + (rank * nsk + col0) * 4 if is_key else off_qs + (rank * nsq + col0) * 4) for b_ in base]

    def v_ptrs(col0):
        return [b_ + off_v + rank * m + col0 for b_ in base]

    if main > 0:
        ops.quant_rows_peers(q[0, give:give + main], main, None, False, world, code_ptrs(False, 0), d, row,
                             scale_ptrs(False, 0), tq)
    gathered.wait()
    km = ops.mean_from_sum(allp[:, 0].sum(dim=0), total, dt)
    vmax = allp[:, 1].amax(dim=0)
    if main > 0:
        ops.quant_rows_peers(k[0, give:give + main], main, km, True, world, code_ptrs(True, 0), d, row,
                             scale_ptrs(True, 0), tk)
        ops.v_convert_peers(v[0, give:give + main], main, vmax, world, v_ptrs(0), d * world * m, world * m, v_scale)
    if edge > 0:
        for i, x in enumerate((q, k, v)):
            if keep > main:
                stage[i, :keep - main].copy_(x[0, give + main:])
    for w in works:
        w.wait()
    if edge > 0:
        ops.quant_rows_peers(stage[0], edge, None, False, world, code_ptrs(False, main), d, row, scale_ptrs(False, mq), tq)
        ops.quant_rows_peers(stage[1], edge, km, True, world, code_ptrs(True, main), d, row, scale_ptrs(True, mk), tk)
        ops.v_convert_peers(stage[2], edge, vmax, world, v_ptrs(main), d * world * m, world * m, v_scale)
    hdl.barrier()      # all stores of all ranks have landed

    qk = hdl.get_buffer(rank, (1, world * m, 2, hp, d), torch.int8, 0)
    v8 = hdl.get_buffer(rank, (hp, d, world * m), torch.uint8, off_v)
    q_scale = hdl.get_buffer(rank, (1, hp, tq), torch.float32, off_qs // 4)
    k_scale = hdl.get_buffer(rank, (1, hp, tk), torch.float32, off_ks // 4)
    if sm_scale is None:
        sm_scale = d ** -0.5
    if peer_out:
        # the way back inside the attention kernel: its TMA-store epilogue writes the rows of destination rank s (this
        # rank's heads) straight into s's [n, H, D] output region over NVLink; a third barrier, one local copy out
        scratch = torch.empty(16 * 136, dtype=torch.uint8, device=dev)     # SAGE_O_PEER_SCRATCH_BYTES
        o_peers = ([b_ + off_o + rank * hp * d * es for b_ in base], n, d, h * d, scratch)
        ops.attention(qk[:, :total, 0], qk[:, :total, 1], v8, q_scale, k_scale,
                      v_scale[rank * hp:(rank + 1) * hp].contiguous(), dt, is_causal, sm_scale, o_peers=o_peers)
        hdl.barrier()  # every rank's epilogue stores have landed
        return hdl.get_buffer(rank, (1, n, h, d), dt, off_o // es).clone()
    o = ops.attention(qk[:, :total, 0], qk[:, :total, 1], v8, q_scale, k_scale,
                      v_scale[rank * hp:(rank + 1) * hp].contiguous(), dt, is_causal, sm_scale)
    r_ = _a2a_heads_out(o, group, world)
    out = torch.empty((1, n, h, d), dtype=dt, device=dev)
    out.view(n, world, hp, d).copy_(r_.transpose(0, 1))
    return out


def ulysses_sageattn(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, group=None,
                     attn_fn: Optional[Callable] = None, head_chunks: int = 1, transport: str = "nccl",
                     comm_sms: int = 0, payload: str = "16bit", _q8_ops=None, **kwargs) -> torch.Tensor:
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
    NCCL's all-to-all kernels (which need SMs) really run next to the attention of another group.
    `payload`: "16bit" ships q, k, v as they are and quantises after the all-to-all; "int8" (B = 1 self-attention with
    head_dim 64 / 128 through the default `sageattn`) quantises the sequence shards first and ships INT8 / e4m3 codes
    and scales - half the inbound bytes (see _ulysses_b1_q8); shapes it does not cover fall back to "16bit".  With
    `transport="p2p"` the quantisers store straight into the peers' symmetric-memory arenas and there is no inbound
    collective at all (_ulysses_b1_q8_p2p)."""
    if kwargs.pop("return_lse", False):
        raise ValueError("ulysses_sageattn does not return the LSE")
    kwargs.pop("tensor_layout", None)
    default_attn = attn_fn is None
    if attn_fn is None:
        from .core import sageattn as attn_fn
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return attn_fn(q, k, v, tensor_layout="NHD", **kwargs)
    h = q.size(2)
    if h % world or k.size(2) % world:
        raise ValueError(f"head counts ({h}, {k.size(2)}) must be divisible by the world size {world}")
    if payload not in ("16bit", "int8"):
        raise ValueError(f"unknown payload {payload!r}")
    if payload == "int8" and q.size(0) == 1 and q.shape == k.shape == v.shape and q.size(3) in (64, 128) and \
            all(t.is_contiguous() for t in (q, k, v)) and q8_geometry(q.size(1), world) is not None and \
            (_q8_ops is not None or (default_attn and q.is_cuda and set(kwargs) <= {"is_causal", "sm_scale", "smooth_k"}
                                     and kwargs.get("smooth_k", True))):
        ops = _q8_ops if _q8_ops is not None else _Q8CudaOps()
        fn = _ulysses_b1_q8_p2p if transport == "p2p" and q.is_cuda and world <= 16 else _ulysses_b1_q8
        return fn(q, k, v, group, world, ops, bool(kwargs.get("is_causal", False)), kwargs.get("sm_scale"))
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
