"""Multi-GPU use of the operator.  The reference itself has no parallel code; its
example gets Ulysses sequence parallelism from xDiT by monkey-patching SDPA
(/root/reference/example/parallel_sageattn_cogvideo.py:8-52).  Every (batch, head) of the
operator is independent, so two layouts cover the named DiT configs:

* head sharding (`shard_heads`): rank r owns heads [r*H/P, (r+1)*H/P) - no communication at all;
* Ulysses (`ulysses_sageattn`): activations arrive sequence-sharded [B, N/P, H, D]; ONE all-to-all
  scatters heads / gathers the sequence for q, k, v together, the operator runs on H/P heads of the
  FULL sequence (so K smoothing, scales and therefore numerics are identical to one GPU), and one
  all-to-all brings O back to [B, N/P, H, D].  NCCL over NVLink/NVSwitch on GPUs; the same code runs
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


def ulysses_sageattn(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, group=None,
                     attn_fn: Optional[Callable] = None, **kwargs) -> torch.Tensor:
    """Sequence-parallel attention.  q, k, v: this rank's sequence block, "NHD" layout
    [B, N/P, H, D] (all ranks hold equally long blocks; pad the sequence to a multiple of P first).
    Returns this rank's block of the output, [B, N/P, H, D].  kwargs go to the attention function
    (default: sageattn); `tensor_layout` is fixed to "NHD" and `return_lse` is not supported here."""
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
    if k.size(2) == h and k.shape == q.shape:
        qkv = _seq_to_head_shard(torch.stack((q, k, v)), group, world)  # one collective for all three
        qh, kh, vh = qkv[0], qkv[1], qkv[2]
    else:  # GQA / cross attention: q and k/v differ in shape
        qh = _seq_to_head_shard(q.unsqueeze(0), group, world)[0]
        kv = _seq_to_head_shard(torch.stack((k, v)), group, world)
        kh, vh = kv[0], kv[1]
    o = attn_fn(qh, kh, vh, tensor_layout="NHD", **kwargs)
    return _head_to_seq_shard(o.contiguous(), group, world)
