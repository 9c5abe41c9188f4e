"""Drop-in for the reference's `lib.mpc_sensitivity` (main/lib/mpc_sensitivity.py), the controller of
main/scenarios/mpc_sensitivity_analysis{,_comulative}.py: identical to lib.mpc except that the cost weights and the
input limits are re-read from '../config/mpc_config_sensitivity.json' on EVERY solve (:150-163), so that the script
can rewrite the file between runs.

    from mpc_for_av_at_intersection_b200.compat.mpc_sensitivity import MPC, MAX_ACCEL

Search order for the file: $MPCB_CONFIG_SENSITIVITY, then the reference's relative path, then the packaged defaults.
The weights go to the device as this problem's weight row (mpcb_session_set_weights_host); the batched form of the
whole sweep is MpcSession.set_weights with one row per (parameter value, vehicle)."""
import json
import os

import numpy as np

from ..config import MpcConfig
from . import mpc as _base
from .mpc import MPCSolutionNotFoundException, smooth_yaw  # noqa: F401


def _find_config():
    env = os.environ.get("MPCB_CONFIG_SENSITIVITY")
    if env:
        return env
    rel = os.path.join("..", "config", "mpc_config_sensitivity.json")
    if os.path.exists(rel):
        return rel
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config", "mpc_config.json")


def _load():
    with open(_find_config(), "r") as f:
        return json.load(f)


_config = _load()
NX = _config["NX"]
NU = _config["NU"]
T = _config["T"]
GOAL_DIS = _config["GOAL_DIS"]
STOP_SPEED = _config["STOP_SPEED"]
MAX_TIME = _config["MAX_TIME"]
MAX_ITER = _config["MAX_ITER"]
DU_TH = _config["DU_TH"]
MAX_DSTEER = np.deg2rad(_config["MAX_DSTEER"])
MAX_ACCEL = _config["MAX_ACCEL"]
MAX_DECEL = _config["MAX_DECEL"]


class MPC(_base.MPC):
    @staticmethod
    def _base_config() -> MpcConfig:
        return MpcConfig.from_dict(dict(_load()))

    def step(self, state):
        cur = MpcConfig.from_dict(dict(_load()))             # mpc_sensitivity.py:150-163, once per solve
        limits = ("MAX_DSTEER", "MAX_ACCEL", "MAX_DECEL")
        if any(getattr(cur, k) != getattr(self._cfg, k) for k in limits):
            for k in limits:                                  # limits live in the handle's config: new session
                setattr(self._cfg, k, getattr(cur, k))
            self._sync_state_to_session()
            self._upload(self._full)
        for k in MpcConfig.WEIGHT_KEYS:
            setattr(self._cfg, k, getattr(cur, k))
        self._ses.set_weights([cur.weight_row()])
        return super().step(state)
