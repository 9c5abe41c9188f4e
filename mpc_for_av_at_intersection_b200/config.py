"""MPC configuration: the JSON schema of the reference (main/config/mpc_config.json, 17 keys) and its
POD mirror for the C ABI (include/mpcb.h: mpcb_config)."""
import json
import os
from dataclasses import dataclass, field
from typing import List

from . import _lib

_DEFAULT_JSON = os.path.join(os.path.dirname(os.path.abspath(__file__)), "config", "mpc_config.json")


@dataclass
class MpcConfig:
    NX: int = 4
    NU: int = 2
    T: int = 13
    w_perp: float = 20.0
    w_para: float = 1.0
    R: List[float] = field(default_factory=lambda: [0.01, 0.01])
    Rd: List[float] = field(default_factory=lambda: [0.01, 1.0])
    Q_v_yaw: List[float] = field(default_factory=lambda: [0.0, 0.5])
    Qf: List[float] = field(default_factory=lambda: [1.0, 1.0, 0.0, 0.5])     # as in the file: unscaled
    GOAL_DIS: float = 1.5
    STOP_SPEED: float = 0.1389
    MAX_TIME: float = 13.0        # unused by the reference too
    MAX_ITER: int = 1
    DU_TH: float = 0.1            # unused by the reference too (early exit is commented out, mpc.py:231-237)
    MAX_DSTEER: float = 30.0      # deg/s as in the file
    MAX_ACCEL: float = 2.0
    MAX_DECEL: float = -10.0
    # not in the JSON: arguments of MPC.__init__ / car dimensions, and solver controls
    dt: float = 0.2
    wheelbase: float = 2.86
    ipm_max_iter: int = 100
    ipm_tol_mu: float = 1e-12
    ipm_tol_p: float = 1e-10
    ipm_tol_d: float = 1e-9
    # inter-vehicle coupling (row A9): car box (car_dimensions.py:82-107) and mpc_intersection.py:81-84 literals
    car_width: float = 2.0
    car_length: float = 3.5
    coupling_horizon: float = 7.0
    frame_window: int = 20
    # variant main/lib/mpc_with_speed.py: speed reference cv (its MAX_SPEED = 25/3.6); 0 = main/lib/mpc.py
    ref_speed: float = 0.0
    # variant main/lib/mpc_jerk.py (NX = 5: acceleration integrated into a fifth state that feeds the speed, jerk penalty)
    jerk_model: bool = False
    jerk_weight: float = 1.0

    @staticmethod
    def from_json(path=None, **overrides):
        with open(path or _DEFAULT_JSON, "r") as f:
            d = json.load(f)
        return MpcConfig.from_dict({**d, **overrides})

    @staticmethod
    def from_dict(d):
        known = set(MpcConfig.__dataclass_fields__)
        unknown = set(d) - known
        if unknown:
            raise KeyError(f"unknown mpc_config keys: {sorted(unknown)}")
        return MpcConfig(**d)

    @staticmethod
    def with_speed(**overrides):
        """The constants main/lib/mpc_with_speed.py:16-36,161,165 hard-codes instead of reading the JSON."""
        d = dict(T=13, w_perp=10.0, w_para=1.0, R=[0.01, 0.01], Rd=[0.01, 1.0], Q_v_yaw=[20.0, 0.5],
                 Qf=[1.0, 1.0, 0.0, 0.5], GOAL_DIS=1.5, STOP_SPEED=0.5 / 3.6, MAX_TIME=13.0, MAX_ITER=1, DU_TH=0.1,
                 MAX_DSTEER=30.0, MAX_ACCEL=2.0, MAX_DECEL=-5.0, ref_speed=25 / 3.6)
        return MpcConfig.from_dict({**d, **overrides})

    @staticmethod
    def jerk(**overrides):
        """The constants main/lib/mpc_jerk.py:16-39,167-171 hard-codes instead of reading the JSON, and its NX = 5 model."""
        d = dict(NX=5, T=13, w_perp=10.0, w_para=1.0, R=[0.01, 0.01], Rd=[0.3, 1.0], Q_v_yaw=[0.0, 0.5],
                 Qf=[1.0, 1.0, 0.0, 0.5], GOAL_DIS=1.5, STOP_SPEED=0.5 / 3.6, MAX_TIME=13.0, MAX_ITER=1, DU_TH=0.1,
                 MAX_DSTEER=30.0, MAX_ACCEL=2.0, MAX_DECEL=-5.0, jerk_model=True, jerk_weight=1.0)
        return MpcConfig.from_dict({**d, **overrides})

    WEIGHT_KEYS = ("w_perp", "w_para", "R", "Rd", "Q_v_yaw", "Qf")

    def weight_row(self, **overrides):
        """[12] row for MpcSession.set_weights: w_perp, w_para, R[2], Rd[2], Q_v_yaw[2], Qf[4] (file units)."""
        g = lambda k: overrides.get(k, getattr(self, k))
        return [float(g("w_perp")), float(g("w_para")), *map(float, g("R")), *map(float, g("Rd")),
                *map(float, g("Q_v_yaw")), *map(float, g("Qf"))]

    def to_c(self):
        c = _lib.mpcb_config()
        _lib.lib().mpcb_default_config(c)
        if self.NU != 2 or self.NX != (5 if self.jerk_model else 4):
            raise ValueError("supported models: NX=4, NU=2 (main/lib/mpc.py) and NX=5 with jerk_model (main/lib/mpc_jerk.py)")
        c.T, c.max_iter = int(self.T), int(self.MAX_ITER)
        c.w_perp, c.w_para = float(self.w_perp), float(self.w_para)
        for i in range(2):
            c.R[i], c.Rd[i], c.Q_v_yaw[i] = float(self.R[i]), float(self.Rd[i]), float(self.Q_v_yaw[i])
        for i in range(4):
            c.Qf[i] = float(self.Qf[i])
        c.goal_dis, c.stop_speed = float(self.GOAL_DIS), float(self.STOP_SPEED)
        c.max_dsteer_deg, c.max_accel, c.max_decel = float(self.MAX_DSTEER), float(self.MAX_ACCEL), float(self.MAX_DECEL)
        c.dt, c.wheelbase = float(self.dt), float(self.wheelbase)
        c.ipm_max_iter = int(self.ipm_max_iter)
        c.ipm_tol_mu, c.ipm_tol_p, c.ipm_tol_d = float(self.ipm_tol_mu), float(self.ipm_tol_p), float(self.ipm_tol_d)
        c.car_width, c.car_length = float(self.car_width), float(self.car_length)
        c.coupling_horizon, c.frame_window = float(self.coupling_horizon), int(self.frame_window)
        c.ref_speed = float(self.ref_speed)
        c.jerk_model, c.jerk_weight = int(bool(self.jerk_model)), float(self.jerk_weight)
        return c
