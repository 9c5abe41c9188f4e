"""Scenario sweeps sharded over ranks (BASELINE.json configs[3]; SURVEY.md section 8d cfg 4, 8e).

A scenario is the atomic unit: every rank owns a contiguous range of scenarios and runs their closed loops with
no collective on the data path.  The only communication is the end-of-run gather of per-scenario results
(torch.distributed: NCCL over NVLink on GPUs, device buffer to device buffer; gloo in the CPU tests).  The reference
has no counterpart -- it iterates vehicles in a Python `for` (main/scenarios/interactive_mpc_temp.py:124) around the
loop body main/scenarios/mpc_intersection.py:95-159."""
import time
from typing import Dict, Tuple

import numpy as np


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split: the first (n_total % world) ranks take one extra scenario.  Returns (first, count)."""
    base, extra = divmod(int(n_total), int(world))
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


def gather_results(local: Dict[str, object], n_total: int, device=None) -> Dict[str, object]:
    """All-gather per-scenario arrays (leading dimension = this rank's scenario count) into rank order.

    torch tensors stay where they are: CUDA tensors are gathered device -> device (NCCL) and come back as CUDA tensors,
    there is no host bounce; numpy arrays (the gloo CPU tests) come back as numpy.  Works without an initialised process
    group (world = 1).  Ragged shards are padded to the largest shard for the collective and trimmed afterwards.  Every
    rank returns the full arrays."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return dict(local)
    world, rank = dist.get_world_size(), dist.get_rank()
    counts = [shard_range(n_total, r, world)[1] for r in range(world)]
    cmax = max(counts)
    out = {}
    for name in sorted(local):
        v = local[name]
        as_numpy = not torch.is_tensor(v)
        t = torch.from_numpy(np.ascontiguousarray(v)) if as_numpy else v.contiguous()
        if device is not None and t.device != torch.device(device):
            t = t.to(device)
        assert t.shape[0] == counts[rank], (name, tuple(t.shape), counts[rank])
        if counts[rank] != cmax:                                   # ragged shard: pad on the tensor's own device
            pad = torch.zeros((cmax,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            pad[:t.shape[0]] = t
            t = pad
        buf = torch.empty((world * cmax,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(buf, t)
        if min(counts) != cmax:
            buf = torch.cat([buf[r * cmax:r * cmax + counts[r]] for r in range(world)], dim=0)
        out[name] = buf.cpu().numpy() if as_numpy else buf
    return out


def checksum(results: Dict[str, object]) -> int:
    """Order-independent-in-nothing, bit-exact 64-bit checksum of a result set (sum of the raw words of every array,
    mod 2^64, arrays in name order, each weighted by its position): equal for two runs iff they agree bit for bit in
    practice; used to show that the union of G shards is the 1-GPU sweep."""
    import torch
    total = 0
    for i, name in enumerate(sorted(results)):
        t = results[name]
        t = torch.from_numpy(np.ascontiguousarray(t)) if not torch.is_tensor(t) else t.contiguous()
        words = t.view(torch.int64) if t.element_size() == 8 else t.view(torch.int32)
        idx = torch.arange(words.numel(), device=words.device, dtype=torch.int64).reshape(words.shape)
        total = (total + (i + 1) * int(((words.to(torch.int64) * (2 * idx + 1)).sum()).item())) % (1 << 64)
    return total


def run_closed_loop_sweep(cfg, courses, n_total: int, nsteps: int, seed: int = 1235, chunk: int = 65536,
                          rank: int = 0, world: int = 1, device: int = 0, keep_history: bool = True, goal_test: bool = True):
    """Closed loops (is_goal -> done | MPC.step | Simulation.step, all on the device, one launch per control step) for
    this rank's shard of a scenario sweep, processed in chunks of at most `chunk` resident scenarios.

    The shard's workload is generated once, up front, on the host (counter-based Philox streams keyed by (seed, block of
    4096 scenario ids): a shard is bit-identical to the same range of the unsharded sweep); everything after that stays on
    the device.  Returns (results, timing): results = CUDA tensors with leading dimension = this rank's scenario count --
    history float32 [count, nsteps, 8] (x y yaw v delta a xref_dev status; NaN rows after the goal), steps_run int32,
    reached_goal int32, final_state float64 [count, 4], fail_count int32, ipm_iters int64 (sum over the steps);
    timing = dict(gen_s, loop_ms (CUDA events around the closed loops only), solves)."""
    import torch
    from .scenarios import make_workload
    from .session import MpcSession
    first, count = shard_range(n_total, rank, world)
    dev = torch.device("cuda", device)
    t0 = time.perf_counter()
    wl = make_workload(courses, count, seed=seed, first=first)
    gen_s = time.perf_counter() - t0
    ses = MpcSession(cfg, courses, min(chunk, max(count, 1)), device=device)
    res = dict(steps_run=torch.zeros(count, dtype=torch.int32, device=dev),
               reached_goal=torch.zeros(count, dtype=torch.int32, device=dev),
               final_state=torch.zeros(count, 4, dtype=torch.float64, device=dev),
               fail_count=torch.zeros(count, dtype=torch.int32, device=dev),
               ipm_iters=torch.zeros(count, dtype=torch.int64, device=dev))
    if keep_history:
        res["history"] = torch.empty(count, nsteps, 8, dtype=torch.float32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    nmax = min(chunk, max(count, 1))
    H = torch.empty(nsteps, nmax, 8, dtype=torch.float64, device=dev)       # one chunk's History block, reused
    loop_ms = 0.0
    for c0 in range(0, count, chunk):
        n = min(chunk, count - c0)
        sl = slice(c0, c0 + n)
        ses.begin(wl["route"][sl], target_ind=wl["target_ind"][sl])
        ses.store("state", np.ascontiguousarray(wl["state"][sl]))
        Hc = H[:, :n] if n == nmax else torch.empty(nsteps, n, 8, dtype=torch.float64, device=dev)
        iters_v = torch.as_tensor(ses.device_view("iters"), device=dev)
        done_v = torch.as_tensor(ses.device_view("done"), device=dev)
        state_v = torch.as_tensor(ses.device_view("state"), device=dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            Hc.fill_(float("nan"))
            e0.record(stream)
            for k in range(nsteps):                                 # one fused launch per control step
                ses.closed_loop_dev(1, Hc[k].data_ptr(), stream.cuda_stream, goal_test=goal_test)
                res["ipm_iters"][sl] += torch.where(done_v == 0, iters_v, torch.zeros_like(iters_v))
            e1.record(stream)
            ran = ~torch.isnan(Hc[:, :, 0])                          # [nsteps, n] rows written = control steps taken
            res["steps_run"][sl] = ran.sum(0).to(torch.int32)
            res["fail_count"][sl] = ((Hc[:, :, 7] >= 2) & ran).sum(0).to(torch.int32)
            res["reached_goal"][sl] = (done_v == 1).to(torch.int32)
            res["final_state"][sl] = state_v
            if keep_history:
                res["history"][sl] = Hc.permute(1, 0, 2)             # float64 [steps, n, 8] -> float32 [n, steps, 8]
        stream.synchronize()
        loop_ms += e0.elapsed_time(e1)
    solves = int(res["steps_run"].sum().item())
    ses.close()
    return res, dict(gen_s=gen_s, loop_ms=loop_ms, solves=solves, first=first, count=count)
