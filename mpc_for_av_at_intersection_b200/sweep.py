"""Scenario sweeps sharded over ranks (BASELINE.json configs[3]).

A scenario is the atomic unit: every rank owns a contiguous range of scenarios and runs their closed loops with
no collective on the data path.  The only communication is the end-of-run gather of per-scenario results
(torch.distributed: NCCL over NVLink on GPUs, gloo in the CPU tests).  The reference has no counterpart -- it
iterates vehicles in a Python `for` (main/scenarios/interactive_mpc_temp.py:124)."""
from typing import Dict, Optional, Tuple

import numpy as np


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split: the first (n_total % world) ranks take one extra scenario.  Returns (first, count)."""
    base, extra = divmod(int(n_total), int(world))
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


def gather_results(local: Dict[str, "np.ndarray"], n_total: int, device=None) -> Optional[Dict[str, np.ndarray]]:
    """All-gather per-scenario arrays (leading dimension = this rank's scenario count) into rank order.

    Works without an initialised process group (world = 1).  Ranks may hold different counts (ragged shards are
    padded to the largest shard for the collective and trimmed afterwards).  Every rank returns the full arrays."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return {k: np.asarray(v) for k, v in local.items()}
    world, rank = dist.get_world_size(), dist.get_rank()
    counts = [shard_range(n_total, r, world)[1] for r in range(world)]
    cmax = max(counts)
    out = {}
    for name in sorted(local):
        v = np.ascontiguousarray(local[name])
        assert v.shape[0] == counts[rank], (name, v.shape, counts[rank])
        t = torch.from_numpy(v)
        pad = torch.zeros((cmax,) + tuple(t.shape[1:]), dtype=t.dtype)
        pad[:t.shape[0]] = t
        if device is not None:
            pad = pad.to(device)
        buf = torch.empty((world * cmax,) + tuple(t.shape[1:]), dtype=t.dtype, device=pad.device)
        dist.all_gather_into_tensor(buf, pad)
        buf = buf.cpu().numpy().reshape((world, cmax) + tuple(t.shape[1:]))
        out[name] = np.concatenate([buf[r, :counts[r]] for r in range(world)], axis=0)
    return out


def run_closed_loop_sweep(cfg, courses, n_total: int, nsteps: int, seed: int = 1235, chunk: int = 65536,
                          rank: int = 0, world: int = 1, device: int = 0, keep_history: bool = False):
    """Closed loops (is_goal -> done | MPC.step | Simulation.step, all on the device) for this rank's shard of a
    scenario sweep, processed in chunks of at most `chunk` scenarios.  Returns per-scenario numpy arrays:
    steps_to_goal (-1 if not reached), final state, fail_count [, history [n, nsteps, 8]]."""
    import torch
    from .scenarios import make_workload
    from .session import MpcSession
    first, count = shard_range(n_total, rank, world)
    ses = MpcSession(cfg, courses, min(chunk, max(count, 1)), device=device)
    dev = torch.device("cuda", device)
    res = dict(steps_to_goal=np.full(count, -1, np.int32), final_state=np.zeros((count, 4)),
               fail_count=np.zeros(count, np.int32))
    if keep_history:
        res["history"] = np.full((count, nsteps, 8), np.nan)
    stream = torch.cuda.Stream(device=dev)
    for c0 in range(0, count, chunk):
        n = min(chunk, count - c0)
        wl = make_workload(courses, n, seed=seed, first=first + c0)
        ses.begin(wl["route"], target_ind=wl["target_ind"])
        ses.store("state", wl["state"])
        with torch.cuda.stream(stream):
            H = torch.full((nsteps, n, 8), float("nan"), dtype=torch.float64, device=dev)
            ses.closed_loop_dev(nsteps, H.data_ptr(), stream.cuda_stream)
            ran = ~torch.isnan(H[:, :, 0])                         # [nsteps, n] rows written = control steps taken
            steps = ran.sum(0).to(torch.int32)
            fails = ((H[:, :, 7] >= 2) & ran).sum(0).to(torch.int32)
        stream.synchronize()
        done = ses.fetch("done")
        sl = slice(c0, c0 + n)
        res["steps_to_goal"][sl] = np.where(done == 1, steps.cpu().numpy(), -1)
        res["fail_count"][sl] = fails.cpu().numpy()
        res["final_state"][sl] = ses.fetch("state")
        if keep_history:
            res["history"][sl] = H.permute(1, 0, 2).cpu().numpy()
    ses.close()
    return res
