"""Result formats: batched device histories in the shapes the reference's analysis code consumes.

The reference keeps one `History` per vehicle (main/lib/simulation.py:64-88: column lists x, y, yaw, v, t, delta, a,
xref_deviation) and plots from it (main/scenarios/mpc_intersection.py:177-271: `plot_trajectories`, `visualize_final`).
The device loops write one row [x, y, yaw, v, delta, a, xref_deviation, status] per vehicle and control step
(include/mpcb.h: mpcb_session_closed_loop_dev).  This module turns a [steps, B, 8] block into per-vehicle `History`
objects, into the exact series each figure of `visualize_final` / `plot_trajectories` draws (as CSV -- there is no
display and no matplotlib on the box; with matplotlib's Agg backend importable the same series are also rendered to
PNG), and into one NPZ for a whole sweep."""
import csv
import os
from typing import List, Optional, Sequence

import numpy as np

from .compat.simulation import History, State

HISTORY_COLUMNS = ("x", "y", "yaw", "v", "t", "delta", "a", "xref_deviation")


def histories_from_device(rows: np.ndarray, state0: np.ndarray, sample_time: float = 0.2) -> List[History]:
    """rows [steps, B, 8] (NaN rows = vehicle already at its goal), state0 [B, 4] -> B `History` objects, each exactly
    what `HistorySimulation` would have accumulated (initial row with a = delta = xref_deviation = 0 included)."""
    rows = np.asarray(rows)
    assert rows.ndim == 3 and rows.shape[2] == 8 and len(state0) == rows.shape[1]
    return [History.from_device_rows(rows[:, b], State(*map(float, state0[b])), sample_time) for b in range(rows.shape[1])]


def history_table(h: History) -> np.ndarray:
    """[n, 8] float64 table in HISTORY_COLUMNS order."""
    return np.array([getattr(h, c) for c in HISTORY_COLUMNS], float).T


def save_history_csv(path: str, h: History):
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(HISTORY_COLUMNS)
        w.writerows(history_table(h).tolist())


def save_sweep_npz(path: str, rows: np.ndarray, state0: np.ndarray, route: Sequence[int], sample_time: float = 0.2,
                   **extra):
    """One file per sweep: the raw [steps, B, 8] block, the initial states, per-vehicle step counts and the
    summary statistics the reference prints per run (steps to goal, mean |xref_deviation|, max |a|)."""
    rows = np.asarray(rows)
    ran = ~np.isnan(rows[:, :, 0])
    steps = ran.sum(0).astype(np.int32)
    with np.errstate(invalid="ignore"):
        dev = np.where(ran, np.abs(rows[:, :, 6]), np.nan)
        mean_dev = np.nanmean(dev, axis=0) if rows.shape[0] else np.zeros(rows.shape[1])
        max_acc = np.nanmax(np.where(ran, np.abs(rows[:, :, 5]), np.nan), axis=0) if rows.shape[0] else np.zeros(rows.shape[1])
    np.savez_compressed(path, history=rows, state0=np.asarray(state0), route=np.asarray(route), steps=steps,
                        mean_xref_deviation=mean_dev, max_abs_accel=max_acc, sample_time=np.array(sample_time),
                        columns=np.array(("x", "y", "yaw", "v", "delta", "a", "xref_deviation", "status")), **extra)


def visualize_final_series(h: History):
    """The three line plots of `visualize_final` (mpc_intersection.py:241-271) as (name, xlabel, ylabel, x, y)."""
    t = np.array(h.t, float)
    return [("speed", "Time [s]", "Speed [km/h]", t, np.array(h.v, float) * 3.6),
            ("acceleration", "Time [s]", "Acceleration [m/s^2]", t, np.array(h.a, float)),
            ("deviation", "Time [s]", "Deviation [m]", t, np.array(h.xref_deviation, float))]


def plot_trajectories_series(ego: History, obstacle_rows: Optional[np.ndarray] = None, dt: float = 0.2):
    """The coloured polylines of `plot_trajectories` (mpc_intersection.py:177-239): per actor (time [s], x, y).
    obstacle_rows: [steps, n_obs, 6] as written by mpcb_session_scripted_loop_dev's obs_history (get() rows)."""
    out = [("ego", np.array(ego.t, float) * dt, np.array(ego.x, float), np.array(ego.y, float))]   # `times * dt`: as written there
    if obstacle_rows is not None:
        obstacle_rows = np.asarray(obstacle_rows)
        for o in range(obstacle_rows.shape[1]):
            ok = ~np.isnan(obstacle_rows[:, o, 0])
            steps = np.flatnonzero(ok)
            out.append((f"obstacle_{o}", steps * dt, obstacle_rows[ok, o, 0], obstacle_rows[ok, o, 1]))
    return out


def export_run(out_dir: str, h: History, obstacle_rows: Optional[np.ndarray] = None, dt: float = 0.2) -> List[str]:
    """Write everything the reference shows at the end of a run into `out_dir`: history.csv, one CSV per figure series
    and, if matplotlib (Agg) is importable, the PNGs.  Returns the files written."""
    os.makedirs(out_dir, exist_ok=True)
    files = [os.path.join(out_dir, "history.csv")]
    save_history_csv(files[0], h)
    for name, xl, yl, x, y in visualize_final_series(h):
        p = os.path.join(out_dir, f"final_{name}.csv")
        np.savetxt(p, np.column_stack([x, y]), delimiter=",", header=f"{xl},{yl}", comments="")
        files.append(p)
    for name, t, x, y in plot_trajectories_series(h, obstacle_rows, dt):
        p = os.path.join(out_dir, f"trajectory_{name}.csv")
        np.savetxt(p, np.column_stack([t, x, y]), delimiter=",", header="time_s,x,y", comments="")
        files.append(p)
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except Exception:
        return files                                       # headless box without matplotlib: the CSVs are the product
    for name, xl, yl, x, y in visualize_final_series(h):
        fig = plt.figure()
        plt.plot(x, y, "-r", label=name)
        plt.grid(True); plt.xlabel(xl); plt.ylabel(yl); plt.tight_layout()
        p = os.path.join(out_dir, f"final_{name}.png")
        fig.savefig(p); plt.close(fig)
        files.append(p)
    fig = plt.figure()
    for name, t, x, y in plot_trajectories_series(h, obstacle_rows, dt):
        plt.scatter(x, y, c=t, s=16 if name == "ego" else 6, cmap="viridis")
    plt.xlim(-40, 40); plt.ylim(-40, 40); plt.xlabel("X"); plt.ylabel("Y")
    p = os.path.join(out_dir, "trajectories.png")
    fig.savefig(p); plt.close(fig)
    files.append(p)
    return files
