"""The N > 1 path on CPU: two processes (gloo), each validating its contiguous shard of one batch with
the CPU checker standing in for the device, then the single exchange step of the path -- the all-gather
of the verdict words -- assembled exactly as bench.py / a multi-GPU host does it over NCCL."""
import os
import socket
import sys

import numpy as np
import pytest

from mr_vamp_b200 import scenes, sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_cover_the_batch_on_word_boundaries():
    for n in [0, 1, 31, 32, 33, 1000, 4097, 10**6 + 7]:
        for world in [1, 2, 3, 4, 8]:
            spans = [sharding.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (lo, hi), (lo2, _) in zip(spans, spans[1:]):
                assert hi == lo2 and lo <= hi
            assert all(lo % 32 == 0 for lo, hi in spans if hi > lo)
            assert all(hi - lo <= 32 * sharding.words_per_rank(n, world) for lo, hi in spans)


def test_unique_id_rendezvous_through_a_file(tmp_path):
    import threading

    path = str(tmp_path / "id")
    uid = bytes(range(128))
    got = {}

    def other(r):
        got[r] = sharding.exchange_unique_id(lambda: b"", r, 3, path, timeout_s=20)

    ts = [threading.Thread(target=other, args=(r,)) for r in (1, 2)]
    for t in ts:
        t.start()
    assert sharding.exchange_unique_id(lambda: uid, 0, 3, path) == uid
    for t in ts:
        t.join()
    assert got == {1: uid, 2: uid}


def test_bit_packing_round_trip():
    rng = np.random.default_rng(0)
    for n in [1, 31, 32, 33, 1000]:
        v = rng.random(n) < 0.5
        w = sharding.pack_bits(v)
        assert w.dtype == np.uint32 and w.size == (n + 31) // 32
        assert np.array_equal(sharding.unpack_bits(w, n), v)
        assert ((int(w[0]) >> 0) & 1) == int(v[0])


def _worker(rank, world, port, n, out_path):
    import torch
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    from oracle import port as oracle_port

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ir = oracle_port.load_ir("panda")
    lo = np.array(ir["scale"]["lo"], np.float32)
    rg = np.array(ir["scale"]["range"], np.float32)
    q = (lo + np.random.default_rng(5).random((n, 7), dtype=np.float32) * rg).astype(np.float32)  # same batch on every rank
    a, b = sharding.shard_range(n, rank, world)
    env = oracle_port.Env(scenes.cage())
    local = oracle_port.fkcc("panda", env, q[a:b]) if b > a else np.zeros(0, bool)

    def allgather(send):  # gloo stands in for the library's ncclAllGather (sharding.NcclComm) on this GPU-less box
        recv = torch.empty(send.size * world, dtype=torch.int32)
        dist.all_gather_into_tensor(recv, torch.from_numpy(send.view(np.int32).copy()))
        return recv.numpy().view(np.uint32)

    full = sharding.gather_verdicts(sharding.pack_bits(local), n, world, allgather)
    if rank == 0:
        np.save(out_path, sharding.unpack_bits(full, n))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [1000, 4097])
def test_two_rank_gather_equals_single_process(tmp_path, n):
    import torch.multiprocessing as mp

    from oracle import port as oracle_port

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "bits.npy")
    mp.spawn(_worker, args=(2, port, n, out), nprocs=2, join=True)
    got = np.load(out)
    ir = oracle_port.load_ir("panda")
    lo = np.array(ir["scale"]["lo"], np.float32)
    rg = np.array(ir["scale"]["range"], np.float32)
    q = (lo + np.random.default_rng(5).random((n, 7), dtype=np.float32) * rg).astype(np.float32)
    want = oracle_port.fkcc("panda", oracle_port.Env(scenes.cage()), q)
    assert np.array_equal(got, want)
