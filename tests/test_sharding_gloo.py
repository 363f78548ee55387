"""CPU tests of the N > 1 path (gloo, world_size 2): every rank derives the same shard plan from the message
sizes without communicating, each message is handled by exactly one rank, and the whole-job throughput is
all ranks' bytes over the slowest rank's time.  (The data path itself has no collective: SURVEY.md §8e.)"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    import tfs_b200 as tfs
    from tfs_b200 import abi, sharding, workloads
    import oracle_bind as ob

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert sharding.rank_world() == (rank, world, rank)
        msgs = workloads.c5_responses(96, seed=5001, max_log2=10)
        sizes = ob.encode_size(tfs.build_batch(msgs, abi.MSG_PREDICT_RESPONSE))
        mine = sharding.my_shard(sizes, rank, world)
        # this rank's share of the work, done by the checker here (no GPU in this container)
        wire = ob.encode_messages([msgs[i] for i in mine], abi.MSG_PREDICT_RESPONSE)
        assert [len(w) for w in wire] == [sizes[i] for i in mine]
        owned = torch.zeros(len(msgs), dtype=torch.int64)
        owned[mine] = 1
        dist.all_reduce(owned)  # test-only check; the product path has no collective
        assert bool((owned == 1).all()), "every message must belong to exactly one rank"
        local_bytes = float(sum(sizes[i] for i in mine))
        total, worst = sharding.aggregate(local_bytes, 10.0 * (rank + 1))
        assert total == float(sum(sizes)) and worst == 10.0 * world
        loads = torch.tensor([local_bytes], dtype=torch.float64)
        gathered = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(gathered, loads)
        spread = max(g.item() for g in gathered) - min(g.item() for g in gathered)
        assert spread <= max(sizes), "shards are balanced to within one message"
        q.put((rank, "ok", sharding.throughput_gbps(total, worst)))
    except Exception as e:  # pragma: no cover
        q.put((rank, f"fail: {e!r}", 0.0))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_ranks_share_a_batch_without_a_data_path_collective():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(30)
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] == "ok" for r in res), res
    assert res[0][2] == res[1][2] > 0


def test_single_process_defaults():
    import tfs_b200 as tfs
    from tfs_b200 import sharding
    assert sharding.aggregate(100.0, 5.0) == (100.0, 5.0)
    assert sharding.my_shard([5, 1, 1, 1], 0, 1) == [0, 1, 2, 3]
    assert sharding.throughput_gbps(2e9, 1000.0) == 2.0
    a, b = sharding.my_shard([8, 1, 1, 1, 1, 1, 1, 1, 1], 0, 2), sharding.my_shard([8, 1, 1, 1, 1, 1, 1, 1, 1], 1, 2)
    assert sorted(a + b) == list(range(9)) and 0 in a and len(b) >= 4
