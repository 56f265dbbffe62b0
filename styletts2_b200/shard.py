"""Batch sharding of ``Decoder.forward`` over the GPUs of one box.

Utterances are independent (InstanceNorm, style and the phase scan are all per item), so the batch
is split into contiguous row ranges, one per rank, and every rank decodes its own rows with its
own handle: no collective on the path.  ``gather=True`` adds one ``all_gather`` of the finished
waveforms for callers that want them everywhere (off the timed path in ``bench.py``).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch


def shard_range(batch: int, world: int, rank: int) -> Tuple[int, int]:
    """rows [lo, hi) of rank ``rank``: contiguous, sizes differ by at most one, empty when world > batch."""
    if batch < 0 or world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad shard request batch={batch} world={world} rank={rank}")
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def decode_sharded(forward: Callable[..., torch.Tensor], asr, F0_curve, N, s, *, world: Optional[int] = None,
                   rank: Optional[int] = None, gather: bool = False, group=None, **kw) -> torch.Tensor:
    """Run ``forward`` (a Decoder module or any callable with its signature) on this rank's rows.

    Inputs are the FULL batch on every rank (or views of it); per-item keyword tensors (``noise``,
    ``har_source``, ``har``) are sliced the same way.  Returns this rank's ``[rows, 1, L]`` waveforms,
    or the full ``[B, 1, L]`` batch on every rank when ``gather`` is set.
    """
    import torch.distributed as dist
    if world is None:
        world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
    B = asr.shape[0]
    if not (F0_curve.shape[0] == N.shape[0] == s.shape[0] == B):
        raise ValueError("asr, F0_curve, N and s must share the batch dimension")
    lo, hi = shard_range(B, world, rank)
    kw = {k: (v[lo:hi] if isinstance(v, torch.Tensor) and v.dim() > 0 and v.shape[0] == B else v) for k, v in kw.items()}
    if hi > lo:
        if getattr(forward, "samples_per_frame", None) is not None:  # a Decoder module: same noise as the full batch
            kw.setdefault("item_offset", lo)
        out = forward(asr[lo:hi], F0_curve[lo:hi], N[lo:hi], s[lo:hi], **kw)
    else:
        out = None
    if not gather or world == 1:
        return out
    # every rank learns the waveform length from rank 0's shard (rank 0 is never empty when B > 0)
    meta = torch.zeros(2, dtype=torch.int64, device=asr.device)
    if out is not None:
        meta[0], meta[1] = out.shape[-1], 1
    dist.all_reduce(meta, op=dist.ReduceOp.MAX, group=group)
    L = int(meta[0])
    sizes = [shard_range(B, world, r) for r in range(world)]
    parts = [torch.empty(h - l, 1, L, dtype=torch.float32, device=asr.device) for l, h in sizes]
    mine = out if out is not None else torch.empty(0, 1, L, dtype=torch.float32, device=asr.device)
    dist.all_gather(parts, mine.contiguous(), group=group) if len({p.shape[0] for p in parts}) == 1 else \
        _all_gather_ragged(parts, mine.contiguous(), rank, group)
    return torch.cat(parts, dim=0)


def _all_gather_ragged(parts, mine, rank, group):
    import torch.distributed as dist
    for r, buf in enumerate(parts):
        if r == rank:
            buf.copy_(mine)
        if buf.numel():
            dist.broadcast(buf, src=r, group=group)
