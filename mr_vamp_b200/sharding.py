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


def gather_verdicts(local_words: np.ndarray, n: int, world: int, allgather) -> np.ndarray:
    """All-gather of the per-rank verdict words.  `local_words`: this rank's uint32 words (any length up to
    words_per_rank); `allgather(send: uint32[w]) -> uint32[w * world]` is the transport in rank order:
    NcclComm.allgather_words below (ncclAllGather over NVLink through the library's vb_comm_*), or whatever the
    host application already has.  Returns the words of all n verdicts in index order."""
    w = words_per_rank(n, world)
    send = np.zeros(w, dtype=np.uint32)
    lw = np.ascontiguousarray(local_words).view(np.uint32)
    send[: lw.size] = lw
    recv = np.asarray(allgather(send)).view(np.uint32)
    assert recv.size == w * world
    return recv[: (n + 31) // 32]


def exchange_unique_id(make_id, rank: int, world: int, path: str, timeout_s: float = 120.0) -> bytes:
    """Rendezvous without a communication library: rank 0 writes the 128-byte NCCL unique id (`make_id()`) to `path`
    (atomically: temporary file + rename), the other ranks wait for it.  One process per GPU on ONE node (the layout
    bench.py is launched in); for several nodes pass the id through the launcher's own store instead."""
    import os
    import time

    if world == 1:
        return make_id()
    if rank == 0:
        uid = make_id()
        tmp = f"{path}.tmp{os.getpid()}"
        with open(tmp, "wb") as f:
            f.write(uid)
        os.replace(tmp, path)
        return uid
    t0 = time.time()
    while time.time() - t0 < timeout_s:
        try:
            with open(path, "rb") as f:
                uid = f.read()
            if len(uid) == 128:
                return uid
        except FileNotFoundError:
            pass
        time.sleep(0.02)
    raise TimeoutError(f"no NCCL unique id at {path} after {timeout_s} s")


class NcclComm:
    """One process per GPU: the library's communicator (vb_comm_init / ncclAllGather over NVLink) with the few
    host-level collectives a launcher needs besides the verdict gather (barrier, max of a float)."""

    def __init__(self, ctx, rank: int, world: int, id_path: str):
        import ctypes as C

        from . import api

        self.ctx, self.rank, self.world = ctx, rank, world
        self._api, self._C = api, C
        self._scratch = None
        if world > 1:
            def make_id():
                buf = (C.c_ubyte * 128)()
                api._check(api.lib().vb_comm_unique_id(buf))
                return bytes(buf)

            uid = exchange_unique_id(make_id, rank, world, id_path)
            raw = (C.c_ubyte * 128).from_buffer_copy(uid)
            api._check(api.lib().vb_comm_init(ctx.handle, world, rank, raw))
            self._send = C.c_void_p()
            self._recv = C.c_void_p()
            api._check(api.lib().vb_dev_alloc(ctx.handle, 0, 4 * 64, C.byref(self._send)))
            api._check(api.lib().vb_dev_alloc(ctx.handle, 0, 4 * 64 * world, C.byref(self._recv)))

    def allgather_small(self, words: np.ndarray) -> np.ndarray:
        """<= 64 uint32 words per rank through device staging buffers -> uint32[world][64]."""
        api, C = self._api, self._C
        send = np.zeros(64, dtype=np.uint32)
        w = np.ascontiguousarray(words).view(np.uint32)
        send[: w.size] = w
        if self.world == 1:
            return send[None, :]
        L, h = api.lib(), self.ctx.handle
        api._check(L.vb_dev_upload(h, 0, self._send, send.ctypes.data_as(C.c_void_p), send.nbytes))
        api._check(L.vb_comm_allgather_bits_d(h, self._send, self._recv, 64))
        out = np.zeros(64 * self.world, dtype=np.uint32)
        api._check(L.vb_dev_download(h, 0, out.ctypes.data_as(C.c_void_p), self._recv, out.nbytes))
        return out.reshape(self.world, 64)

    def barrier(self):
        self.ctx.sync()
        self.allgather_small(np.zeros(1, np.uint32))

    def max_float(self, v: float) -> float:
        got = self.allgather_small(np.array([v], dtype=np.float32))
        return float(got[:, 0].view(np.float32).max())

    def close(self):
        if self.world > 1:
            self._api.lib().vb_comm_destroy(self.ctx.handle)
