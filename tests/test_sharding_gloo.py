"""The N>1 path on CPU: world_size-2 gloo processes run the sequence sharding, the barrier / max-over-ranks timing
reduction and the record gather that bench.py uses under torchrun (no data-path collective exists)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from isosurfacesuperresolution_b200 import sharding, synthetic


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world_size, port, n_seq, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        r, w = sharding.world()
        start, stop = sharding.sequence_range(n_seq, r, w)
        # every rank generates only ITS sequences; seeds are per sequence so the union is shard-count independent
        mine = [synthetic.synthetic_gbuffer(1, 8, 8, 1000 + s, T=2)[0] for s in range(start, stop)]
        checksum = float(sum(np.float64(m).sum() for m in mine))
        sharding.barrier()
        ms = sharding.max_over_ranks(10.0 + 5.0 * r)          # the slowest rank defines the step time
        total = sharding.sum_over_ranks(stop - start)
        recs = sharding.gather_records({"rank": r, "units": (stop - start) * 100, "ms": 10.0 + 5.0 * r,
                                        "range": (start, stop), "checksum": checksum})
        q.put((r, ms, total, recs, sharding.weak_range(16, r), sharding.rank_seed(4321, r)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_seq", [5, 16])
def test_two_rank_sequence_sharding(n_seq):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_seq, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    single = float(sum(np.float64(synthetic.synthetic_gbuffer(1, 8, 8, 1000 + s, T=2)[0]).sum() for s in range(n_seq)))
    for rank, ms, total, recs, weak, seed in results:
        assert ms == 15.0 and total == n_seq
        assert [r["rank"] for r in recs] == [0, 1]
        ranges = [r["range"] for r in recs]
        assert ranges[0][0] == 0 and ranges[0][1] == ranges[1][0] and ranges[1][1] == n_seq      # disjoint cover
        assert abs(sum(r["checksum"] for r in recs) - single) < 1e-6 * max(1.0, abs(single))     # same data as 1 rank
        assert sharding.aggregate_throughput(recs, unit_scale=1.0) == pytest.approx(n_seq * 100 / 15e-3)
        assert weak == (16 * rank, 16 * rank + 16) and seed == 4321 + rank


def test_sequence_range_properties():
    for n in (0, 1, 7, 64, 513):
        for w in (1, 2, 3, 8):
            rs = [sharding.sequence_range(n, r, w) for r in range(w)]
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            sizes = [b - a for a, b in rs]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.sequence_range(4, 2, 2)
    assert sharding.world() == (0, 1) and sharding.gather_records({"a": 1}) == [{"a": 1}]
    assert sharding.max_over_ranks(3.5) == 3.5
