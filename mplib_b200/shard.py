"""Sharding of a configuration batch (or of planning queries) over the GPUs of one box.

Every configuration is independent and the compiled scene is replicated on each GPU, so the path needs no data-path
collective: rank r owns a contiguous slice of the batch and the only inter-GPU traffic is ONE all-gather of
the packed verdict words after the last kernel (N/8 bytes in total; NCCL over NVLink when the tensors are CUDA
tensors, gloo for the CPU tests).  One process per GPU (``torchrun``); ``torch.distributed`` is plumbing only.
"""
from __future__ import annotations

import numpy as np


def shard_bounds(n: int, world: int):
    """Slice boundaries [b_0 .. b_world]: equal contiguous shards, aligned to 32 configurations so that every packed
    verdict word belongs to exactly one rank (rank r owns [b_r, b_{r+1}))."""
    words = (n + 31) // 32
    b = [min(n, 32 * ((words * r) // world)) for r in range(world)] + [n]
    return b


def shard_range(n: int, rank: int, world: int):
    b = shard_bounds(n, world)
    return b[rank], b[rank + 1]


def pack_bits(flags) -> np.ndarray:
    """bool[n] -> uint32 words, bit (i & 31) of word i >> 5 (the layout of ``mpb_collide_batch``)."""
    flags = np.asarray(flags, dtype=bool)
    pad = (-flags.shape[0]) % 32
    if pad:
        flags = np.concatenate([flags, np.zeros(pad, bool)])
    return np.packbits(flags, bitorder="little").view(np.uint32).copy()


def unpack_bits(words, n: int) -> np.ndarray:
    return np.unpackbits(np.asarray(words, dtype=np.uint32).view(np.uint8), bitorder="little")[:n].astype(bool)


def gather_verdict_words(local_words, n: int, group=None):
    """All-gather of the per-rank verdict words -> the words of the whole batch, on every rank.

    ``local_words``: torch int32 tensor with the words of this rank's shard (CUDA tensor under NCCL, CPU under gloo).
    Shards are word-aligned (``shard_bounds``), so the result is a plain concatenation."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    b = shard_bounds(n, world)
    counts = [(b[r + 1] - b[r] + 31) // 32 for r in range(world)]
    m = max(counts) if counts else 0
    buf = torch.zeros(m, dtype=torch.int32, device=local_words.device)
    buf[: local_words.numel()] = local_words
    out = torch.empty(world * m, dtype=torch.int32, device=local_words.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    parts = [out[r * m: r * m + counts[r]] for r in range(world)]
    return torch.cat(parts) if parts else out


def collide_sharded(collide_words_fn, states, group=None):
    """Run ``collide_words_fn(states_local) -> int32 word tensor`` on this rank's shard and gather the verdicts.

    ``states``: the full [n, D] batch (every rank holds or can generate it); returns bool[n] on every rank."""
    import torch.distributed as dist

    n = states.shape[0]
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    lo, hi = shard_range(n, rank, world)
    words = collide_words_fn(states[lo:hi])
    allw = gather_verdict_words(words, n, group)
    return unpack_bits(allw.cpu().numpy().view(np.uint32), n)


def shard_queries(n_queries: int, rank: int, world: int):
    """Planning queries: contiguous equal shards (1024 / G per GPU in configs[3])."""
    return (n_queries * rank) // world, (n_queries * (rank + 1)) // world


def gather_objects(local, group=None):
    """Gather per-rank Python results (statuses, ragged paths) -- the planners' final result gather."""
    import torch.distributed as dist

    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, local, group=group)
    return out
