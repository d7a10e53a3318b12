"""CPU, world_size 2, gloo: the N>1 path.  Requests shard by id (id % world) with no data-path
collective; each rank routes its shard independently (here with the CPU oracle standing in for
its GPU), and the merged records equal the unsharded result.  Also covers the max-over-ranks
timing reduction bench.py uses."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from polykey_service_b200.shard import merge_shards, select_requests, shard_ids


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pk_oracle as O
        table = O.default_table()
        stream = O.gen_batch(O.SEED_BASE + 4, total, 0, 3000, mode=-1)       # same seeded stream on every rank
        mine = select_requests(stream, shard_ids(total, world, rank))
        rc, rec = O.route_batch(table, mine, threads=2)
        assert rc == 0
        # bench.py's reduction: every rank times its own work, the job time is the max
        t = torch.tensor([float(rank + 1) * 1.5], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == world * 1.5
        gathered = [None] * world
        dist.all_gather_object(gathered, rec.tobytes())
        if rank == 0:
            recs = [np.frombuffer(b, dtype=O.DECISION_DTYPE) for b in gathered]
            merged = merge_shards(total, world, recs)
            rc, whole = O.route_batch(table, stream, threads=2)
            q.put(bool(rc == 0 and (merged == whole).all()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_sharded_stream_equals_unsharded(world, oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 1001, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_shard_helpers_roundtrip(oracle):
    b = oracle.gen_batch(3, 257, 0, 500, mode=0)
    for world in (1, 2, 4, 8):
        ids = [shard_ids(257, world, r) for r in range(world)]
        assert sorted(np.concatenate(ids).tolist()) == list(range(257))
        t = oracle.default_table()
        whole = oracle.route_batch(t, b)[1]
        recs = [oracle.route_batch(t, select_requests(b, i))[1] for i in ids]
        assert (merge_shards(257, world, recs) == whole).all()


def test_numa_binding_helpers(tmp_path):
    """sysfs parsing and the affinity choice of shard.bind_to_gpu_numa (fake sysfs tree; the
    process's real affinity is restored)."""
    import os
    from polykey_service_b200 import shard
    assert shard.parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert shard.parse_cpulist("") == []
    dev = tmp_path / "0000:1b:00.0"
    dev.mkdir()
    allowed = sorted(os.sched_getaffinity(0))
    (dev / "local_cpulist").write_text(f"{allowed[0]}\n")
    assert shard.gpu_local_cpus("00000000:1B:00.0", str(tmp_path)) == [allowed[0]]
    assert shard.gpu_local_cpus("0000:ff:00.0", str(tmp_path)) == []
    try:
        if len(allowed) > 1:
            assert shard.bind_to_gpu_numa("0000:1b:00.0", str(tmp_path)) == [allowed[0]]
            assert os.sched_getaffinity(0) == {allowed[0]}
        os.sched_setaffinity(0, allowed)
        assert shard.bind_to_gpu_numa("0000:ff:00.0", str(tmp_path)) == []          # unknown device
        (dev / "local_cpulist").write_text("100000-100001\n")                        # outside the cpuset
        assert shard.bind_to_gpu_numa("0000:1b:00.0", str(tmp_path)) == []
        assert os.sched_getaffinity(0) == set(allowed)
    finally:
        os.sched_setaffinity(0, allowed)
