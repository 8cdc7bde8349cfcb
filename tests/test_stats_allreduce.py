"""N>1 host logic on CPU: env sharding + the single statistics all-reduce over gloo, world size 2."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dexrobot_isaac_b200 import stats as S


def test_shard_range_partitions_exactly():
    for total in (1, 7, 64, 4096, 131072, 131075):
        for world in (1, 2, 3, 8):
            spans = [S.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (a, n), (b, _) in zip(spans, spans[1:]):
                assert a + n == b
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        S.shard_range(10, 2, 2)


def test_pack_unpack_roundtrip():
    rows = 5
    totals = torch.arange(S.N_COUNTERS + 1, dtype=torch.int64)
    st = torch.arange(4 * (2 + rows), dtype=torch.float64).reshape(4, 2 + rows) + 1
    out = S.unpack(S.pack(totals, st), [f"c{i}" for i in range(rows)])
    assert out["steps"] == S.N_COUNTERS and out["success"] == S.C["DEX_CNT_SUCCESS"]
    assert out["episodes"]["all"]["n"] == 1.0 and out["episodes"]["all"]["components"]["c0"] == 3.0
    assert "episodes" not in S.unpack(S.pack(totals))
    with pytest.raises(ValueError):
        S.pack(torch.zeros(3, dtype=torch.int64))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rows = 3
    totals = torch.full((S.N_COUNTERS + 1,), rank + 1, dtype=torch.int64)
    st = torch.full((4, 2 + rows), float(10 * (rank + 1)), dtype=torch.float64)
    payload = S.pack(totals, st)
    S.all_reduce(payload)
    start, count = S.shard_range(131072, rank, world)
    q.put((rank, payload.tolist(), start, count))
    dist.destroy_process_group()


def test_gloo_world2_allreduce():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, v0, s0, c0), (r1, v1, s1, c1) = got
    assert v0 == v1  # every rank holds the global sum
    assert v0[0] == 3.0 and v0[S.N_COUNTERS + 1] == 30.0
    assert (s0, c0, s1, c1) == (0, 65536, 65536, 65536)
