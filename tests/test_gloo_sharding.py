"""Multi-rank host logic on CPU: world_size-2 (and 3, ragged) gloo process groups exercise the shard arithmetic and
the end-of-run gather that the GPU sweep uses over NCCL.  No CUDA here."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_total, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mpc_for_av_at_intersection_b200 import CourseSet, load_intersection_routes
        from mpc_for_av_at_intersection_b200.scenarios import make_workload
        from mpc_for_av_at_intersection_b200.sweep import gather_results, shard_range
        cs = CourseSet(load_intersection_routes())
        first, count = shard_range(n_total, rank, world)
        wl = make_workload(cs, count, seed=9, first=first)
        # stand-in for per-scenario results of this shard: something every rank can recompute for the check
        local = dict(final_state=wl["state"] * 2.0, steps_to_goal=(wl["route"] * 3 + 1).astype(np.int32),
                     ids=np.arange(first, first + count, dtype=np.int64))
        full = gather_results(local, n_total)
        import torch
        tens = gather_results({"ids": torch.from_numpy(local["ids"])}, n_total)      # tensors in -> tensors out (the GPU path)
        assert torch.is_tensor(tens["ids"]) and torch.equal(tens["ids"], torch.arange(n_total))
        ref = make_workload(cs, n_total, seed=9)
        ok = (np.array_equal(full["ids"], np.arange(n_total)) and
              np.array_equal(full["final_state"], ref["state"] * 2.0) and
              np.array_equal(full["steps_to_goal"], (ref["route"] * 3 + 1).astype(np.int32)))
        q.put((rank, bool(ok), int(count)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total", [(2, 1000), (3, 1001)])
def test_gather_over_gloo(world, n_total):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in got)
    assert sum(c for _, _, c in got) == n_total


def test_shard_range_partitions_exactly():
    from mpc_for_av_at_intersection_b200.sweep import shard_range
    for n in (0, 1, 7, 4096, 1048576, 1000003):
        for w in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == n
            for (f0, c0), (f1, _) in zip(spans, spans[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
