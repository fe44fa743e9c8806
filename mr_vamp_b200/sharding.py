"""Host-side sharding of query batches over one process per GPU.

Every configuration and every edge is independent and the environment is replicated, so a batch is
split into contiguous index ranges whose boundaries fall on multiples of 32 (verdict words are never
shared between ranks) and the only exchange is an all-gather of the verdict words.  Inside one
process the C library does the same split over the devices of a context (vb_capi.cu: shards_of).
"""
from __future__ import annotations

import numpy as np


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """[lo, hi) of `rank`: ceil(n / world) rounded up to a multiple of 32, clipped to n."""
    per = (((n + world - 1) // world) + 31) // 32 * 32
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def words_per_rank(n: int, world: int) -> int:
    """Verdict words every rank contributes to the all-gather (ranks with a short shard pad with zeros)."""
    per = (((n + world - 1) // world) + 31) // 32 * 32
    return per // 32


def pack_bits(valid: np.ndarray) -> np.ndarray:
    """bool[n] -> uint32 words, bit i%32 of word i//32 (the layout of vb_fkcc's valid_bits)."""
    v = np.asarray(valid, dtype=bool)
    pad = (-v.size) % 32
    if pad:
        v = np.concatenate([v, np.zeros(pad, bool)])
    return np.packbits(v, bitorder="little").view(np.uint32)


def unpack_bits(words: np.ndarray, n: int) -> np.ndarray:
    return np.unpackbits(np.ascontiguousarray(words, dtype=np.uint32).view(np.uint8), bitorder="little")[:n].astype(bool)


def gather_verdicts(local_words, n: int, group=None):
    """All-gather of the per-rank verdict words with torch.distributed (NCCL for CUDA tensors over
    NVLink, gloo for CPU tensors in the tests).  `local_words`: int32/uint32 torch tensor with this
    rank's words (any length up to words_per_rank).  Returns the words of all n verdicts in index order."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    w = words_per_rank(n, world)
    send = torch.zeros(w, dtype=torch.int32, device=local_words.device)
    send[: local_words.numel()] = local_words.view(torch.int32)
    recv = torch.empty(w * world, dtype=torch.int32, device=local_words.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    return recv[: (n + 31) // 32]
