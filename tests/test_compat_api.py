"""Host logic of the reference-compatible faces (no GPU): module constants follow main/lib/mpc.py:17-36, the
scalar Simulation/History reproduce the reference's plant (checked against the golden histories), CarDimensions
numbers follow main/lib/car_dimensions.py."""
import os

import numpy as np

from conftest import golden


def test_car_dimensions():
    from mpc_for_av_at_intersection_b200.compat.car_dimensions import BicycleModelDimensions, PriusDimensions
    b = BicycleModelDimensions()
    assert b.distance_back_to_front_wheel == 2.86 and b.bounding_box_size == (2.0, 3.5)
    assert b.radius == 2.0 / 2 ** .5 and b.center_point_offset == (1.43, 0.0)
    assert np.allclose(b.circle_centers, [[1.43 + 0.75, 0.0], [1.43 - 0.75, 0.0]])
    assert BicycleModelDimensions(skip_back_circle_collision_checking=True).circle_centers.shape == (1, 2)
    p = PriusDimensions()
    assert p.distance_back_to_front_wheel == 4 and p.bounding_box_size == (2.04, 4.84)
    assert PriusDimensions(0.5).distance_back_to_front_wheel == 2.0


def test_scalar_plant_reproduces_reference_history():
    from mpc_for_av_at_intersection_b200.compat.car_dimensions import BicycleModelDimensions
    from mpc_for_av_at_intersection_b200.compat.simulation import HistorySimulation, State
    g = golden("closed_loop_T13.npz")
    off = g["history_off"]
    for i in (0, 5, 12):
        h = g["history"][off[i]:off[i + 1]]
        sim = HistorySimulation(BicycleModelDimensions(), 0.2, State(*h[0, :4]))
        for k in range(1, len(h)):
            s = sim.step(a=h[k, 5], delta=h[k, 4], xref_deviation=h[k, 6])
            assert (s.x, s.y, s.yaw, s.v) == tuple(h[k, :4]), "the plant is bit-exact float64"
        hist = sim.history
        assert len(hist.x) == len(h) and np.allclose(hist.t, 0.2 * np.arange(1, len(h) + 1))
        assert np.array_equal(np.array(hist.delta), h[:, 4]) and np.array_equal(np.array(hist.a), h[:, 5])


def test_history_from_device_rows():
    from mpc_for_av_at_intersection_b200.compat.simulation import History, State
    rows = np.full((5, 8), np.nan)
    rows[:3] = np.arange(24).reshape(3, 8)
    h = History.from_device_rows(rows, State(1, 2, 3, 4), 0.2)
    assert h.x == [1, 0.0, 8.0, 16.0] and h.a == [0.0, 5.0, 13.0, 21.0] and h.delta == [0.0, 4.0, 12.0, 20.0]


def test_module_constants_follow_reference_rules():
    import importlib
    import json
    import os
    import mpc_for_av_at_intersection_b200.compat.mpc as m
    cfg = json.load(open(m.CONFIG_PATH))
    assert m.T == cfg["T"] and m.MAX_ACCEL == cfg["MAX_ACCEL"] and m.MAX_DECEL == cfg["MAX_DECEL"]
    assert np.array_equal(m.Qf, np.diag(cfg["Qf"]) * cfg["T"])           # mpc.py:25
    assert m.MAX_DSTEER == np.deg2rad(cfg["MAX_DSTEER"])                 # mpc.py:34
    assert m.R.shape == (2, 2) and m.Q_v_yaw.shape == (2, 2)
    for name in ("MPC", "MPCSolutionNotFoundException", "smooth_yaw", "NX", "NU", "w_perp", "w_para", "Rd",
                 "GOAL_DIS", "STOP_SPEED", "MAX_TIME", "MAX_ITER", "DU_TH"):
        assert hasattr(m, name)
    # $MPCB_CONFIG overrides, as a scratch '../config/mpc_config.json' would for the reference
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        cfg20 = dict(cfg, T=20)
        path = os.path.join(d, "c.json")
        json.dump(cfg20, open(path, "w"))
        os.environ["MPCB_CONFIG"] = path
        try:
            m2 = importlib.reload(m)
            assert m2.T == 20 and m2.Qf[0, 0] == cfg["Qf"][0] * 20
        finally:
            del os.environ["MPCB_CONFIG"]
            importlib.reload(m)


def test_smooth_yaw_in_place_on_strided_view():
    import mpc_oracle as mo
    from mpc_for_av_at_intersection_b200.compat.mpc import smooth_yaw
    rng = np.random.default_rng(1)
    traj = rng.normal(size=(300, 3))
    traj[:, 2] = (np.cumsum(rng.normal(0, 0.9, 300)) + np.pi) % (2 * np.pi) - np.pi
    want = mo.smooth_yaw(traj[:, 2].copy())
    out = smooth_yaw(traj[:, 2])           # a strided view, as main/scenarios/mpc_intersection.py:71 passes
    assert np.array_equal(traj[:, 2], want) and out is not None


def test_config_roundtrip_and_unknown_keys():
    import pytest
    from mpc_for_av_at_intersection_b200 import MpcConfig
    c = MpcConfig.from_json(T=20)
    cc = c.to_c()
    assert cc.T == 20 and cc.max_dsteer_deg == 30.0 and list(cc.Qf) == [1.0, 1.0, 0.0, 0.5]
    with pytest.raises(KeyError):
        MpcConfig.from_dict({"T": 13, "bogus": 1})


def test_workload_shards_compose():
    from mpc_for_av_at_intersection_b200 import CourseSet, load_intersection_routes
    from mpc_for_av_at_intersection_b200.scenarios import make_workload
    cs = CourseSet(load_intersection_routes())
    full = make_workload(cs, 10000, seed=5)
    parts = [make_workload(cs, 2500, seed=5, first=2500 * r) for r in range(4)]
    for k in ("route", "state", "target_ind"):
        assert np.array_equal(full[k], np.concatenate([p[k] for p in parts]))
    odd = make_workload(cs, 777, seed=5, first=4000)
    assert np.array_equal(odd["state"], full["state"][4000:4777])
    assert full["state"][:, 3].min() >= 0 and full["state"][:, 3].max() <= 30 / 3.6
    assert set(np.unique(full["route"])) == set(range(12))


def test_flop_model_matches_survey_table():
    import bench
    # F at K=15, Nc=660
    for T, want in ((10, 0.11e6), (13, 0.20e6), (20, 0.60e6), (40, 3.95e6), (80, 28.7e6)):
        got = bench.algorithmic_flops(T, 660, 15)
        assert abs(got - want) / want < 0.06, (T, got, want)


def test_result_formats_from_device_rows(tmp_path):
    """a [steps, B, 8] device block -> per-vehicle History identical to the one HistorySimulation builds
    when fed the same steps; CSV / NPZ exports and the figure series of visualize_final / plot_trajectories."""
    from conftest import golden
    from mpc_for_av_at_intersection_b200 import results
    from mpc_for_av_at_intersection_b200.compat.simulation import History
    g = golden("closed_loop_T13.npz")
    off = g["history_off"]
    loops = [0, 12]
    n = [int(off[i + 1] - off[i]) - 1 for i in loops]
    K = max(n) + 2
    rows = np.full((K, 2, 8), np.nan)
    for j, i in enumerate(loops):
        ref = g["history"][off[i]:off[i + 1]]                 # x y yaw v delta a dev, row 0 = initial state
        rows[:n[j], j, :7] = ref[1:]
        rows[:n[j], j, 7] = 0
    state0 = np.stack([g["history"][off[i], :4] for i in loops])
    hs = results.histories_from_device(rows, state0, 0.2)
    for j, i in enumerate(loops):
        ref = g["history"][off[i]:off[i + 1]]
        tab = results.history_table(hs[j])
        assert tab.shape == (n[j] + 1, 8)
        assert np.array_equal(tab[:, [0, 1, 2, 3, 5, 6, 7]], ref)
        assert np.allclose(tab[:, 4], 0.2 * np.arange(1, n[j] + 2))      # History.store adds dt before storing (simulation.py:77-78)
    sel = np.flatnonzero(g["loop"] == 12)
    files = results.export_run(str(tmp_path / "run"), hs[1], g["obs"][sel], 0.2)
    assert len(files) >= 7 and all(os.path.getsize(f) > 0 for f in files)
    sp = np.loadtxt(tmp_path / "run" / "final_speed.csv", delimiter=",", skiprows=1)
    assert np.allclose(sp[:, 1], np.array(hs[1].v) * 3.6)
    tr = np.loadtxt(tmp_path / "run" / "trajectory_obstacle_1.csv", delimiter=",", skiprows=1)
    assert tr.shape == (len(sel), 3) and tr[0, 1] == 30.0
    results.save_sweep_npz(str(tmp_path / "sweep.npz"), rows, state0, [0, 9])
    z = np.load(tmp_path / "sweep.npz")
    assert list(z["steps"]) == n and z["history"].shape == rows.shape and z["mean_xref_deviation"].shape == (2,)
