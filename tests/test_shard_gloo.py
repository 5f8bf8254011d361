"""world_size-2 (and 3) gloo tests of the multi-GPU host logic: contiguous word-aligned shards, one
all-gather of packed verdict words, query sharding with an object gather.  The per-rank checker here is the CPU
oracle (tests may use it); on the GPU box the same functions wrap ``mpb_collide_batch`` (bench.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mplib_b200 import shard as S


def test_shard_bounds_properties():
    for n in (0, 1, 31, 32, 33, 1000, 1 << 20, (1 << 20) + 5):
        for w in (1, 2, 3, 4, 8):
            b = S.shard_bounds(n, w)
            assert b[0] == 0 and b[-1] == n and all(x <= y for x, y in zip(b, b[1:]))
            assert all(x % 32 == 0 for x in b[:-1])
            sizes = np.diff(b)
            if n >= 32 * w:
                assert sizes.max() - sizes.min() <= 32 + (n % 32)
    v = np.random.default_rng(0).random(1000) < 0.3
    assert np.array_equal(S.unpack_bits(S.pack_bits(v), 1000), v)
    assert S.shard_queries(1024, 3, 8) == (384, 512)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    from tests.common import compile_desc, make_panda_world, random_q

    desc = compile_desc(make_panda_world())
    q = random_q(n, 61)  # every rank generates the same seeded batch and checks only its shard
    calls = []

    def collide_words(q_local):
        calls.append(q_local.shape[0])
        v, _ = O.collide_batch(desc, q_local.astype(np.float32).astype(np.float64))
        return torch.from_numpy(S.pack_bits(v).view(np.int32).copy())

    verdict = S.collide_sharded(collide_words, q)
    lo, hi = S.shard_range(n, rank, world)
    assert calls == [hi - lo]
    # planning-style object gather
    qlo, qhi = S.shard_queries(10, rank, world)
    gathered = S.gather_objects([(i, f"path{i}") for i in range(qlo, qhi)])
    flat = [x for part in gathered for x in part]
    assert [i for i, _ in flat] == list(range(10))
    np.save(os.path.join(out_dir, f"verdict_{rank}.npy"), verdict)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 1000), (2, 4096), (3, 777)])
def test_collide_sharded_gloo(tmp_path, world, n):
    from oracle import oracle as O
    from tests.common import compile_desc, make_panda_world, random_q

    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    desc = compile_desc(make_panda_world())
    q = random_q(n, 61)
    ref, _ = O.collide_batch(desc, q.astype(np.float32).astype(np.float64))
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), f"verdict_{r}.npy"))
        assert np.array_equal(got, ref)
    assert ref.any() and not ref.all()
