"""BASELINE.json configs[4]: batch x horizon sweep of the fused step (latency at small B, throughput at large B).
Runs bench.py once per cell and writes one JSON table.  usage: python scripts/sweep_config5.py out.json"""
import json, os, subprocess, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/config5_sweep.json"
rows = []
for T in (10, 13, 20, 40, 80):
    for B in (1, 8, 64, 4096, 65536):
        if T == 80 and B == 65536:
            steps = 3
        elif T >= 40 and B >= 4096:
            steps = 6
        else:
            steps = 20
        cmd = [sys.executable, os.path.join(REPO, "bench.py"), "--no-cpu-baseline", "--steps", str(steps), "--warmup", "3",
               "--batch", str(B), "--horizon", str(T)]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        try:
            d = json.loads(r.stdout.strip().splitlines()[-1])
            rows.append(dict(T=T, B=B, steps=steps, solves_per_s=d["value"], ms_per_step=d["ms_per_step"], p50_step_ms=d["p50_step_ms"],
                             e2e_solves_per_s=d["e2e"]["value"], frac=d["roofline"]["frac"], mean_ipm_iters=d["config"]["mean_ipm_iters"],
                             failed_solves=d["config"]["failed_solves"], sm_mhz=d["clocks"]["sm_mhz"]))
            print(rows[-1], flush=True)
        except Exception as e:
            rows.append(dict(T=T, B=B, error=str(e), stderr=r.stderr[-500:]))
            print("FAILED", T, B, r.stderr[-300:], flush=True)
json.dump(dict(config="BASELINE.json configs[4]: cfg-2 generator, closed loop, fused step + plant, L2 flushed between steps", rows=rows),
          open(out, "w"), indent=1)
