#!/usr/bin/env python
"""bench.py -- MPC solves/sec of the batched B200 MPC tracking loop (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B] [--horizon T]

Workload (config.workload): BASELINE.json configs[1] -- 4,096 independent ego vehicles per GPU on randomised
intersection scenarios, horizon N=20, closed loop (every step: MPC.step for the whole batch, then the plant
step).  One "step" = one control step of the whole batch = B solves.  N>1: one process per GPU (torchrun),
scenarios sharded by rank, no data-path collective ("scaling": "weak"); the only collective is the timing
reduction (max over ranks).

  value     solves/s with states resident in HBM (one fused kernel launch per control step, CUDA-event timed
            per step on the launching stream, L2 flushed between timed steps)
  e2e       solves/s through the host-buffer C-ABI call mpcb_session_step_host (MpcSession.step): vehicle
            states copied host->device and applied controls + status device->host inside the timed region
  roofline  FP64-FMA roofline of the step kernel: algorithmic flops F(T, Nc, K) of SURVEY.md section 8d x
            solves per launch / average launch duration, against the FP64 FMA peak measured live
  cpu_baseline  the oracle port (numpy float64, oracle/mpc_oracle.py) on the host cores, bounded sample

`--impl reference` times that CPU port alone, on all host cores, and prints the same line shape."""
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


def cpu_arm(args, courses, cfg, per_core_problems, steps_each, cores=None):
    """Time the CPU port (oracle/mpc_port.cpp, all host cores) on a bounded sample of the bench workload.
    Returns (dict, solves/s)."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import cpu_baseline
    from mpc_for_av_at_intersection_b200.scenarios import make_workload
    cores = cores or os.cpu_count() or 1
    n = min(args.batch, 4096)
    wl = make_workload(courses, n, seed=args.seed, first=0)
    cpu_baseline.run_cpp(courses, wl, min(n, 64), 1, cfg, cores=cores)          # load / page in
    r = cpu_baseline.run_cpp(courses, wl, n, steps_each, cfg, cores=cores, min_seconds=args.cpu_seconds)
    v = r["solves"] / r["seconds"]
    return ({"value": v, "unit": UNIT, "cores": r["cores"], "kind": "port",
             "sample": f"first {n} problems of the workload x {r['steps']} closed-loop steps = {r['solves']} solves "
                       f"in {r['seconds']:.1f} s wall on {r['cores']} threads; plain C++ float64 port oracle/mpc_port.cpp "
                       f"(generic dense condensing + dense Mehrotra IPM in place of cvxpy+ECOS, which like "
                       f"/root/reference is not on this box)"}, v)


def run_reference(args):
    """--impl reference: the CPU port on all host cores; a step = one control step of a bounded sample (<= 4096
    problems) of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from mpc_for_av_at_intersection_b200 import CourseSet, MpcConfig, load_intersection_routes
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import cpu_baseline
    from mpc_for_av_at_intersection_b200.scenarios import make_workload
    cfg = MpcConfig.from_json(T=args.horizon)
    courses = CourseSet(load_intersection_routes())
    cores = os.cpu_count() or 1
    per_step = min(args.batch, 4096) if cfg.T <= 20 else min(args.batch, 256)
    wl = make_workload(courses, per_step, seed=args.seed, first=0)
    if args.warmup > 0:
        cpu_baseline.run_cpp(courses, wl, per_step, min(args.warmup, 3), cfg, cores=cores)
    r = cpu_baseline.run_cpp(courses, wl, per_step, args.steps, cfg, cores=cores)
    v = r["solves"] / r["seconds"]
    sample = (f"{per_step} problems x {args.steps} closed-loop steps = {r['solves']} solves in {r['seconds']:.1f} s wall on "
              f"{r['cores']} threads; plain C++ float64 port of main/lib/mpc.py (oracle/mpc_port.cpp: generic dense condensing + "
              f"dense Mehrotra IPM; the reference's cvxpy+ECOS stack is not installable offline), {r['failed']} failed solves")
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * r["seconds"] / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"BASELINE.json configs[1]: independent ego vehicles on randomised intersection "
                               f"scenarios, horizon N={cfg.T}, closed loop; bounded sample of {per_step} problems "
                               f"per step", "horizon": cfg.T, "problems_per_step": per_step, "seed": args.seed},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample},
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
        cpu, _ = cpu_arm(args, courses, cfg, args.cpu_problems_per_core, args.cpu_steps)

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
    ses = MpcSession(cfg, courses, B, device=local).begin(wl["route"], target_ind=wl["target_ind"])
    ses.store("state", wl["state"])
    stream = torch.cuda.Stream(device=dev)
    sptr = stream.cuda_stream
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    # ---- device-resident closed loop: value -----------------------------------------------------------------
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            ses.closed_loop_dev(1, 0, sptr, goal_test=False)
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
    l0 = ses.launches
    t_wall0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for k in range(args.steps):
            flush.zero_()                                   # L2 flush, outside the event pair
            scan_acc += (nc_v - tgt_v).clamp_(min=0)        # course points scanned by this step's lookup
            ev[k][0].record(stream)
            ses.closed_loop_dev(1, 0, sptr, goal_test=False)
            ev[k][1].record(stream)
            iters_acc += iters_v
            bad_acc += (status_v >= 2).sum()
    barrier()
    t_wall1 = time.perf_counter()
    launches = ses.launches - l0
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

    # ---- end to end through the host-buffer C ABI -----------------------------------------------------------------
    ses.begin(wl["route"], target_ind=wl["target_ind"])
    state = np.ascontiguousarray(wl["state"])
    L = cfg.wheelbase
    for _ in range(args.warmup):
        d, a = ses.step(state)
        state = host_plant_step(state, d, a, L, cfg.dt)
    barrier()
    e2e_s = 0.0
    for _ in range(args.steps):
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
            "config": {"workload": f"BASELINE.json configs[1]: {B} independent ego vehicles per GPU on randomised "
                                   f"intersection scenarios (12 A* routes), horizon N={T}, closed loop "
                                   f"(MPC.step + plant step per control step)",
                       "batch_per_gpu": B, "horizon": T, "seed": args.seed, "sharding": f"scenarios x{world}",
                       "l2": "flushed between timed steps (256 MiB memset outside the event pairs)",
                       "mean_ipm_iters": k_mean, "mean_course_points_scanned": scan_mean,
                       "failed_solves": n_bad},
            "clocks": clocks,
            "e2e": {"value": B * world * args.steps / e2e_s_max, "unit": UNIT,
                    "h2d_bytes_per_step": 32 * B, "d2h_bytes_per_step": 20 * B,
                    "api": "MpcSession.step -> mpcb_session_step_host (host buffers, pinned staging inside)"},
            "gpu_launches": launches,
            "roofline": {"bound": "fp64_fma", "achieved": ach_tflops, "peak": peak, "unit": "TFLOP/s",
                         "frac": ach_tflops / peak if peak else None, "traffic": ncu_traffic(B, T),
                         "kernel": (f"mpcb::mpc_step_kernel<{T},1,8> (one warp per problem, one warp per CTA, 8 CTAs/SM)" if T <= 31
                                    else f"mpcb::mpc_step_long_kernel<128> (one CTA per problem, T={T})"), "flops_per_solve": F,
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
    ap.add_argument("--cpu-problems-per-core", type=int, default=4)
    ap.add_argument("--cpu-steps", type=int, default=12, help="closed-loop steps of the cpu_baseline sample (<= 4096 problems)")
    ap.add_argument("--ref-problems-per-core", type=int, default=2)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="minimum wall time of the cpu_baseline sample")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    return run_reference(args) if args.impl == "reference" else run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
