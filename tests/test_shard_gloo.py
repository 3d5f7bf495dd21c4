"""The N>1 path on CPU: world_size-2 gloo processes shard a run by batch, each produces the counters of its own
batches (with the oracle standing in for the GPU -- this tests the sharding and the collective, not the kernels),
and the all-reduced counters equal the single-process run."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
TOTAL, BATCH, SEED = 5000, 700, 0x47415352


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _counters_for(ol, schedule):
    c = np.zeros(4, np.int64)
    for first, n in schedule:
        b = ol.generate(ol.gen_config(SEED, 100, 0.10), first, n)
        g = ol.oracle_greedy(b, ol.make_params())
        nw = ol.oracle_nw(b, 1, 1, 1)
        c += np.asarray(ol.oracle_metrics(b, g, nw), np.int64)
    return c


def _worker(rank, world_size, port, out_dir):
    sys.path.insert(0, HERE)
    import conftest
    import oracle_lib as ol
    pkg = conftest.load_pkg()
    from gasma_b200 import shard
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        assert shard.world()[:2] == (rank, world_size)
        sched = shard.batch_schedule(TOTAL, BATCH, rank, world_size)
        mine = torch.from_numpy(_counters_for(ol, sched))
        shard.reduce_counters(mine)
        slow = shard.max_over_ranks(1.0 + rank)
        np.save(os.path.join(out_dir, "r%d.npy" % rank), np.concatenate([mine.numpy(), [int(slow)], [len(sched)]]))
    finally:
        dist.destroy_process_group()


def test_schedule_covers_every_pair_once(pkg):
    from gasma_b200 import shard
    for total, batch, ws in [(0, 10, 2), (1, 10, 4), (5000, 700, 2), (10**6, 10**6, 8), (10**6 + 1, 10**5, 3)]:
        seen = []
        for r in range(ws):
            seen += shard.batch_schedule(total, batch, r, ws)
        seen.sort()
        pos = 0
        for first, n in seen:
            assert first == pos and 0 < n <= batch
            pos += n
        assert pos == total
        edges = [shard.shard_range(total, r, ws) for r in range(ws)]
        assert edges[0][0] == 0 and edges[-1][1] == total and all(a[1] == b[0] for a, b in zip(edges, edges[1:]))


@pytest.mark.timeout(300)
def test_two_rank_gloo_run_equals_single_process(pkg, ol, tmp_path):
    from gasma_b200 import shard
    single = _counters_for(ol, shard.batch_schedule(TOTAL, BATCH, 0, 1))
    assert single[0] == TOTAL
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        got = np.load(os.path.join(tmp_path, "r%d.npy" % r))
        assert got[:4].tolist() == single.tolist()      # every rank holds the global sum
        assert got[4] == 2                               # max over ranks of (1 + rank)
        assert got[5] == (4 if r == 0 else 4)            # 8 batches of <=700 pairs dealt round-robin
