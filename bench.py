#!/usr/bin/env python
"""bench.py -- MPC solves/sec of the batched B200 MPC tracking loop (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B] [--horizon T]
                    [--workload config2|sweep]

Workload (config.workload): BASELINE.json configs[1] -- 4,096 independent ego vehicles per GPU on randomised
intersection scenarios, horizon N=20, closed loop (every step: MPC.step for the whole batch, then the plant
step).  One "step" = one control step of the whole batch = B solves.  N>1: one process per GPU (torchrun),
scenarios sharded by rank, no data-path collective ("scaling": "weak"); the only collective is the timing
reduction (max over ranks).

The timed steps are ALWAYS control steps 5..24 of the closed loops that start at the workload's cold start
(PHASE_LEAD untimed lead-in steps, then up to PHASE_LEN timed ones; for --steps > 20 the same start is re-seeded and
the lead-in repeated, untimed).  The interior-point iterations per solve depend on how far a loop has settled, so a
number that followed --steps/--warmup could not be compared between rounds; this one does not.  `mean_ipm_iters` and
`ipm_iterations_per_sec` (= value x mean_ipm_iters) are printed so that two kernels can be compared even if the
iteration count moves.

  value     solves/s with states resident in HBM (one fused kernel launch per control step, CUDA-event timed
            per step on the launching stream, L2 flushed between timed steps)
  e2e       solves/s through the host-buffer C-ABI call mpcb_session_step_host (MpcSession.step): vehicle
            states copied host->device and applied controls + status device->host inside the timed region
  roofline  FP64-FMA roofline of the step kernel: algorithmic flops F(T, Nc, K) of SURVEY.md section 8d x
            solves per launch / average launch duration, against the FP64 FMA peak measured live
  cpu_baseline  the plain C++ port of main/lib/mpc.py (oracle/mpc_port.cpp, -O3 -march=native, all host threads) on a
            bounded sample of the same workload and the same control steps; only the solver call is inside its clock

`--impl reference` times that CPU port alone, on all host cores, and prints the same line shape.
`--workload sweep` runs BASELINE.json configs[3] (1,048,576-scenario closed-loop sweep, strong scaling, NCCL gather)."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

METRIC = "mpc_solves_per_sec"
UNIT = "solves/s"
PHASE_LEAD, PHASE_LEN = 5, 20      # timed steps = control steps 5..24 of the closed loops, whatever --steps/--warmup


def algorithmic_flops(T, nc_scanned, K):
    """F(T, Nc, K), FMA = 2 flops, n = 2T, m = 8T."""
    n, m = 2 * T, 8 * T
    condense = 6.0 * nc_scanned + 40.0 * T + 32.0 * T * (T + 1) + (32.0 / 3.0) * T ** 3 + 16.0 * T ** 2
    per_iter = n ** 3 / 3.0 + 6.0 * n ** 2 + 10.0 * m
    return condense + K * per_iter


def algorithmic_bytes(T, nc_scanned):
    """HBM bytes per solve."""
    return 16.0 * nc_scanned + 8.0 * (4 + 2 * T + 3 * (T + 1) + 2 * T + 4 * (T + 1)) + 16.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md, clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        rows = [ln for (t, ln) in self.lines if (t0 is None or t >= t0 - 0.05) and (t1 is None or t <= t1 + 0.15)]
        if not rows:
            rows = [ln for (_, ln) in self.lines]
        sm, smax, reasons, power = [], [], set(), []
        for ln in rows:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def step_kernel_name(T):
    """The kernel mpcb_api.cu's dispatch_step launches for this horizon (default environment)."""
    if T <= 15 and os.environ.get("MPCB_HALF_WARP", "1")[:1] != "0":
        return f"mpcb::mpc_step_half_kernel<{T},1,8> (two problems per warp, one per half-warp; one warp per CTA, 8 CTAs/SM)"
    if T <= 31:
        return f"mpcb::mpc_step_kernel<{T},1,8> (one warp per problem, one warp per CTA, 8 CTAs/SM)"
    nt = 256 if T > 40 else 64
    return f"mpcb::mpc_step_long_kernel<{nt},{1 if T > 40 else 4},false> (one CTA of {nt} threads per problem, {1 if T > 40 else 4} CTAs/SM, longest-first launch order)"


def workload_string(B, T):
    """config.workload, identical in both arms (the driver compares the strings)."""
    return (f"BASELINE.json configs[1]: {B} independent ego vehicles per GPU on randomised intersection scenarios "
            f"(12 A* routes), horizon N={T}, closed loop (MPC.step + plant step per control step), timed steps = "
            f"control steps {PHASE_LEAD}..{PHASE_LEAD + PHASE_LEN - 1} from the cold start")


def host_cpu():
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import mpc_port
    return mpc_port.cpu_model()[0]


def probe_cvxpy_ecos():
    """BASELINE.md section 3.1 / SURVEY.md section 8c: is the reference's real solver stack on this box?"""
    try:
        import cvxpy  # noqa: F401
        import ecos  # noqa: F401
        return os.path.basename(os.path.dirname(getattr(cvxpy, "__file__", "") or "")) != "cvxpy_shim" and hasattr(cvxpy, "ECOS")
    except Exception:
        return False


def cpu_sample(args, courses, cfg, n, steps, min_seconds, cores=None):
    """Time the CPU port (oracle/mpc_port.cpp, all host cores) on the first `n` problems of the bench workload, same
    control steps as the GPU arm.  Returns (cpu_baseline dict, run dict)."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import cpu_baseline
    from mpc_for_av_at_intersection_b200.scenarios import make_workload
    cores = cores or os.cpu_count() or 1
    wl = make_workload(courses, n, seed=args.seed, first=0)
    cpu_baseline.run_cpp(courses, wl, min(n, 64), 1, cfg, cores=cores, lead_in=0)          # load / page in
    r = cpu_baseline.run_cpp(courses, wl, n, steps, cfg, cores=cores, min_seconds=min_seconds, lead_in=PHASE_LEAD, phase_len=PHASE_LEN)
    v = r["solves"] / r["seconds"]
    return ({"value": v, "unit": UNIT, "cores": r["cores"], "kind": "port", "cpu_model": host_cpu(),
             "sample": f"first {n} problems of the workload x {r['steps']} timed control steps (steps {PHASE_LEAD}..{PHASE_LEAD + PHASE_LEN - 1} "
                       f"of the closed loops, re-seeded) = {r['solves']} solves in {r['seconds']:.1f} s of solver time on {r['cores']} threads "
                       f"({r['failed']} failed); plain C++ float64 port oracle/mpc_port.cpp (-O3 -march=native; generic dense condensing + "
                       f"dense Mehrotra IPM in place of cvxpy+ECOS, which like /root/reference is not on this box)"}, r)


def run_reference(args):
    """--impl reference: the CPU port on all host cores; a step = one control step of a bounded sample (<= 4096
    problems) of the same workload, same control steps as the GPU arm.  This process never maps libmpcb.so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, load_intersection_routes
    cfg = MpcConfig.from_json(T=args.horizon)
    courses = CourseSet(load_intersection_routes(), smooth="numpy")
    per_step = min(args.batch, 4096) if cfg.T <= 20 else min(args.batch, 256)
    cpu, r = cpu_sample(args, courses, cfg, per_step, args.steps, 0.0)
    v = cpu["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * r["seconds"] / max(r["steps"], 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_string(args.batch, cfg.T), "batch_per_gpu": args.batch, "horizon": cfg.T, "seed": args.seed,
                   "sample": f"bounded sample: the first {per_step} problems of the workload per step", "failed_solves": r["failed"]},
        "cpu_baseline": cpu, "cvxpy_ecos_present": probe_cvxpy_ecos(),
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def ncu_traffic(B, T):
    """dram__bytes_read.sum + dram__bytes_write.sum of one step-kernel launch, from the committed ncu --set full
    capture of the same configuration (profiles/*_step_traffic.json); None when no capture matches."""
    import glob
    for f in sorted(glob.glob(os.path.join(REPO, "profiles", "*_step_traffic.json")), reverse=True):
        try:
            d = json.load(open(f))
            if d.get("batch") == B and d.get("horizon") == T:
                return d["dram_bytes_read"] + d["dram_bytes_write"]
        except Exception:
            pass
    return None


def run_b200(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, MpcSession, _lib, load_intersection_routes
    from mpc_for_av_at_intersection_b200.scenarios import host_plant_step, make_workload
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    cfg = MpcConfig.from_json(T=args.horizon)
    courses = CourseSet(load_intersection_routes())
    T, B = cfg.T, args.batch

    # ---- CPU baseline first (fork pool before CUDA is initialised), rank 0 at N=1 only ---------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu, _ = cpu_sample(args, courses, cfg, min(args.batch, 4096) if cfg.T <= 20 else min(args.batch, 256), args.cpu_steps,
                            args.cpu_seconds)

    import numpy as np
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    lib = _lib.lib()

    wl = make_workload(courses, B, seed=args.seed, first=rank * B)     # shard = contiguous scenario range
    ses = MpcSession(cfg, courses, B, device=local)
    stream = torch.cuda.Stream(device=dev)
    sptr = stream.cuda_stream
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def reseed_dev(lead):
        """the workload's cold start, then `lead` untimed control steps on the device"""
        ses.begin(wl["route"], target_ind=wl["target_ind"])
        ses.store("state", wl["state"])
        with torch.cuda.stream(stream):
            for _ in range(lead):
                ses.closed_loop_dev(1, 0, sptr, goal_test=False)
        torch.cuda.synchronize(dev)

    # ---- device-resident closed loop: value -----------------------------------------------------------------
    reseed_dev(args.warmup)                                 # W untimed warm-up steps (clocks, caches, function attributes)
    reseed_dev(PHASE_LEAD)
    barrier()
    sampler = ClockSampler(local).start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    iters_acc = torch.zeros(B, dtype=torch.int64, device=dev)
    scan_acc = torch.zeros(B, dtype=torch.int64, device=dev)
    iters_v = torch.as_tensor(ses.device_view("iters"), device=dev)
    tgt_v = torch.as_tensor(ses.device_view("target_ind"), device=dev)
    nc_v = torch.as_tensor(ses.device_view("ncourse"), device=dev)
    status_v = torch.as_tensor(ses.device_view("status"), device=dev)
    bad_acc = torch.zeros((), dtype=torch.int64, device=dev)
    exc_acc = torch.zeros((), dtype=torch.int64, device=dev)      # status 3: the reference's nearest-index Exception, not a failed solve
    l0 = ses.launches
    lead_launches = 0
    t_wall0 = time.perf_counter()
    for k in range(args.steps):
        if k and k % PHASE_LEN == 0:                        # same loops again: the timed steps stay control steps 5..24
            reseed_dev(PHASE_LEAD)
            lead_launches += PHASE_LEAD
        with torch.cuda.stream(stream):
            flush.zero_()                                   # L2 flush, outside the event pair
            scan_acc += (nc_v - tgt_v).clamp_(min=0)        # course points scanned by this step's lookup
            ev[k][0].record(stream)
            ses.closed_loop_dev(1, 0, sptr, goal_test=False)
            ev[k][1].record(stream)
            iters_acc += iters_v
            bad_acc += (status_v == 2).sum()
            exc_acc += (status_v == 3).sum()
    barrier()
    t_wall1 = time.perf_counter()
    launches = ses.launches - l0 - lead_launches
    step_ms = [a.elapsed_time(b) for a, b in ev]
    dev_ms = float(sum(step_ms))
    clocks = sampler.stop(t_wall0, t_wall1)
    tmax = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dev_ms_max = float(tmax.item())
    k_mean = float(iters_acc.double().mean().item()) / max(args.steps, 1)
    scan_mean = float(scan_acc.double().mean().item()) / max(args.steps, 1)
    n_bad = int(bad_acc.item())
    n_exc = int(exc_acc.item())

    # ---- end to end through the host-buffer C ABI, same control steps ------------------------------------------------
    L = cfg.wheelbase

    def reseed_host(lead):
        ses.begin(wl["route"], target_ind=wl["target_ind"])
        st = np.ascontiguousarray(wl["state"])
        for _ in range(lead):
            d, a = ses.step(st)
            st = host_plant_step(st, d, a, L, cfg.dt)
        return st

    state = reseed_host(PHASE_LEAD)
    barrier()
    e2e_s = 0.0
    for k in range(args.steps):
        if k and k % PHASE_LEN == 0:
            state = reseed_host(PHASE_LEAD)
        t0 = time.perf_counter()
        d, a = ses.step(state)                               # H2D state, kernel, D2H controls + status, sync
        e2e_s += time.perf_counter() - t0
        state = host_plant_step(state, d, a, L, cfg.dt)      # the caller's simulation, not timed
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s_max = float(te.item())

    # ---- roofline --------------------------------------------------------------------------------------------------
    peak = C_double_peak(lib, local)
    F = algorithmic_flops(T, scan_mean, k_mean)
    avg_launch_s = dev_ms * 1e-3 / max(launches, 1)
    ach_tflops = F * B / avg_launch_s / 1e12
    hbm_peak = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(REPO, "MEASURED_PEAKS.json")) else 6650.0
    ach_gbs = algorithmic_bytes(T, scan_mean) * B / avg_launch_s / 1e9

    if rank == 0:
        value = B * world * args.steps / (dev_ms_max * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms_max / args.steps, "p50_step_ms": sorted(step_ms)[len(step_ms) // 2],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_string(B, T),
                       "batch_per_gpu": B, "horizon": T, "seed": args.seed, "sharding": f"scenarios x{world}",
                       "l2": "flushed between timed steps (256 MiB memset outside the event pairs)",
                       "mean_ipm_iters": k_mean, "ipm_iterations_per_sec": value * k_mean,
                       "mean_course_points_scanned": scan_mean, "failed_solves": n_bad, "nearest_index_exceptions": n_exc},
            "cvxpy_ecos_present": probe_cvxpy_ecos(),
            "clocks": clocks,
            "e2e": {"value": B * world * args.steps / e2e_s_max, "unit": UNIT,
                    "h2d_bytes_per_step": 32 * B, "d2h_bytes_per_step": 20 * B,
                    "api": "MpcSession.step -> mpcb_session_step_host (host buffers, pinned staging inside)"},
            "gpu_launches": launches,
            "roofline": {"bound": "fp64_fma", "achieved": ach_tflops, "peak": peak, "unit": "TFLOP/s",
                         "frac": ach_tflops / peak if peak else None, "traffic": ncu_traffic(B, T),
                         "kernel": step_kernel_name(T), "flops_per_solve": F,
                         "peak_source": "measured live (mpcb_measure_fp64_peak: register-resident DFMA chains)",
                         "hbm": {"achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                                 "peak_source": "MEASURED_PEAKS.json hbm_gbs" if os.path.exists(
                                     os.path.join(REPO, "MEASURED_PEAKS.json")) else "fallback"}},
            "wall_s_timed_region": t_wall1 - t_wall0,
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_sweep(args):
    """--workload sweep: BASELINE.json configs[3] -- `--scenarios` single-ego closed loops (the config-2 generator), STRONG
    scaling: rank r owns shard_range(n, r, N), chunks of `--chunk` resident scenarios, `--sweep-steps` control steps with
    the goal test on, float32 history [scenarios/N, steps, 8] and per-scenario statistics kept on the device, then ONE
    NCCL all-gather device -> device per array and one D2H of the union on rank 0.  The workload is generated on the host
    before the clock starts (reported as gen_s)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, _lib, load_intersection_routes
    from mpc_for_av_at_intersection_b200.sweep import checksum, gather_results, run_closed_loop_sweep
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    cfg = MpcConfig.from_json(T=args.horizon)
    courses = CourseSet(load_intersection_routes())
    n, K = args.scenarios, args.sweep_steps

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    # warm-up: a small sweep (function attributes, clocks, NCCL communicator)
    w, _ = run_closed_loop_sweep(cfg, courses, 4096 * world, max(args.warmup, 3), seed=args.seed, chunk=args.chunk, rank=rank,
                                 world=world, device=local)
    gather_results({"steps_run": w["steps_run"]}, 4096 * world)
    if world > 1:
        # NCCL sets up its large-message path on the first big collective (one run measured 39 ms for the 2.7 GB gather at N = 2
        # against 4.4 ms in the others): a 64 MB-per-rank all-gather outside the clock, so that the timed one is steady state
        import torch.distributed as dist
        big = torch.zeros(16 << 20, dtype=torch.float32, device=dev)
        out = torch.empty(world * big.numel(), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(out, big)
        torch.cuda.synchronize(dev)
        del big, out
    del w
    barrier()
    host = None
    if rank == 0:                                               # pinned landing buffers of the union, allocated before the clock
        shapes = dict(history=((n, K, 8), torch.float32), steps_run=((n,), torch.int32), reached_goal=((n,), torch.int32),
                      final_state=((n, 4), torch.float64), fail_count=((n,), torch.int32), ipm_iters=((n,), torch.int64))
        host = {k: torch.empty(sh, dtype=dt, pin_memory=True) for k, (sh, dt) in shapes.items()}
    sampler = ClockSampler(local).start()
    t0 = time.perf_counter()
    res, tm = run_closed_loop_sweep(cfg, courses, n, K, seed=args.seed, chunk=args.chunk, rank=rank, world=world, device=local)
    barrier()
    t_loop = time.perf_counter()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    full = gather_results(res, n)
    g1.record()
    barrier()
    gather_ms = g0.elapsed_time(g1)
    t_gather = time.perf_counter()
    gathered_bytes = sum(v.numel() * v.element_size() for v in full.values())
    if rank == 0:                                               # one D2H of the union
        for k, v in full.items():
            host[k].copy_(v, non_blocking=True)
        torch.cuda.synchronize(dev)
    t_end = time.perf_counter()
    clocks = sampler.stop(t0, t_end)
    stats = torch.tensor([tm["loop_ms"], tm["gen_s"], gather_ms], dtype=torch.float64, device=dev)
    solves = torch.tensor([tm["solves"]], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(solves, op=dist.ReduceOp.SUM)
    loop_ms, gen_s, gather_ms = (float(x) for x in stats.tolist())
    if rank == 0:
        total = int(solves.item())
        csum = checksum(full)
        k_mean = float(full["ipm_iters"].sum().item()) / max(total, 1)
        line = {
            "metric": METRIC, "value": total / (loop_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": loop_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"BASELINE.json configs[3]: {n}-scenario closed-loop sweep (config-2 generator, seed {args.seed}), "
                                   f"horizon N={cfg.T}, {K} control steps with the goal test on, sharded by scenario over {world} GPU(s) "
                                   f"in chunks of {args.chunk}, NCCL all-gather of history + statistics at the end",
                       "scenarios": n, "horizon": cfg.T, "seed": args.seed, "chunk": args.chunk, "mean_ipm_iters": k_mean,
                       "l2": "inputs larger than L2 (65536 resident problems x 25.6 KB of per-problem state)"},
            "clocks": clocks,
            "sweep": {"solves": total, "reached_goal": int(full["reached_goal"].sum().item()), "failed_solves": int(full["fail_count"].sum().item()),
                      "failed_solves_counts": "status >= 2: failed solves and nearest-index Exceptions of the reference (scripts/triage_sweep_failures.py tells them apart)",
                      "workload_gen_s_untimed": gen_s, "closed_loops_ms_device_max_over_ranks": loop_ms,
                      "wall_s_loops": t_loop - t0, "wall_s_gather": t_gather - t_loop, "wall_s_d2h_rank0": t_end - t_gather,
                      "wall_s_total": t_end - t0, "gather_ms": gather_ms, "gathered_bytes": gathered_bytes,
                      "gather_gbs_per_gpu_received": gathered_bytes * (world - 1) / world / (gather_ms * 1e-3) / 1e9 if world > 1 else None,
                      "d2h_bytes": gathered_bytes, "checksum": f"{csum:016x}",
                      "history": f"float32 [{n}, {K}, 8] on the device, union on every rank after the gather"},
            "e2e": {"value": total / (t_end - t0), "unit": UNIT, "h2d_bytes_per_step": 44 * n // K // world, "d2h_bytes_per_step": gathered_bytes // K,
                    "api": "sweep.run_closed_loop_sweep + gather_results: host workload in, gathered union on the host out (wall clock incl. "
                           "workload upload, closed loops, NCCL gather, D2H)"},
            "gpu_launches": K * ((n // world + args.chunk - 1) // args.chunk),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def C_double_peak(lib, device):
    import ctypes as C
    v = C.c_double(0.0)
    rc = lib.mpcb_measure_fp64_peak(device, C.byref(v))
    return float(v.value) if rc == 0 else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="problems per GPU")
    ap.add_argument("--horizon", type=int, default=20)
    ap.add_argument("--seed", type=int, default=1235)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-steps", type=int, default=12, help="closed-loop steps of the cpu_baseline sample (<= 4096 problems)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="minimum solver time of the cpu_baseline sample")
    ap.add_argument("--workload", default="config2", choices=["config2", "sweep"], help="sweep = BASELINE.json configs[3]")
    ap.add_argument("--scenarios", type=int, default=1 << 20, help="sweep: scenarios over ALL ranks (strong scaling)")
    ap.add_argument("--sweep-steps", type=int, default=80, help="sweep: closed-loop control steps per scenario")
    ap.add_argument("--chunk", type=int, default=65536, help="sweep: scenarios resident per launch")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_sweep(args) if args.workload == "sweep" else run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
