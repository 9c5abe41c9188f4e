"""Plant-side types with the reference's names (main/lib/simulation.py:12-88, main/bicycle/main.py:28-41).

The scalar `Simulation` here is host arithmetic for one vehicle (a scenario script's own plant, exactly the
reference's float64 operation order); the batched plant is the CUDA kernel behind mpcb_plant_step /
mpcb_session_closed_loop_dev."""
import math
from dataclasses import dataclass
from typing import List, Optional

import numpy as np


@dataclass
class State:
    x: float = 0.0
    y: float = 0.0
    yaw: float = 0.0
    v: float = 0.0


class Simulation:
    MAX_STEER = np.deg2rad(45.0)
    MAX_SPEED = 30.0 / 3.6
    MIN_SPEED = -5.

    def __init__(self, car_dimensions, sample_time: float, initial_state: State):
        self._L = car_dimensions.distance_back_to_front_wheel
        self._dt = sample_time
        self._s = State(initial_state.x, initial_state.y, initial_state.yaw, initial_state.v)

    def step(self, a: float, delta: float) -> State:
        s, dt = self._s, self._dt
        delta = max(min(delta, Simulation.MAX_STEER), -Simulation.MAX_STEER)
        # Bicycle.step: all right-hand sides use the pre-update heading and speed
        xdot, ydot = s.v * np.cos(s.yaw), s.v * np.sin(s.yaw)
        yawdot = (s.v / self._L) * np.tan(delta)
        v = max(min(s.v + a * dt, Simulation.MAX_SPEED), Simulation.MIN_SPEED)
        self._s = State(s.x + xdot * dt, s.y + ydot * dt, s.yaw + yawdot * dt, v)
        return State(self._s.x, self._s.y, self._s.yaw, self._s.v)


class History:
    """Column lists x, y, yaw, v, t, delta, a, xref_deviation (simulation.py:64-88)."""

    def __init__(self, sample_time: float):
        self.x: List[float] = []
        self.y: List[float] = []
        self.yaw: List[float] = []
        self.v: List[float] = []
        self.t: List[float] = []
        self.delta: List[float] = []
        self.a: List[float] = []
        self.xref_deviation: List[float] = []
        self._sample_time = sample_time

    def get_current_time(self) -> float:
        return self.t[-1] if self.t else 0.

    def store(self, state: State, a: float, delta: float, xref_deviation: Optional[float] = None):
        now = self.get_current_time() + self._sample_time
        for col, val in ((self.x, state.x), (self.y, state.y), (self.yaw, state.yaw), (self.v, state.v), (self.t, now),
                         (self.delta, delta), (self.a, a),
                         (self.xref_deviation, np.nan if xref_deviation is None else xref_deviation)):
            col.append(val)

    @staticmethod
    def from_device_rows(rows: np.ndarray, initial_state: State, sample_time: float) -> "History":
        """Build a History from one problem's rows [n,8] of mpcb_session_closed_loop_dev's history buffer
        (x, y, yaw, v, delta, a, xref_deviation, status); NaN rows (finished problem) are dropped."""
        h = History(sample_time)
        h.store(initial_state, a=0., delta=0., xref_deviation=0.)
        for r in rows:
            if math.isnan(r[0]):
                continue
            h.store(State(r[0], r[1], r[2], r[3]), a=r[5], delta=r[4], xref_deviation=r[6])
        return h


class HistorySimulation(Simulation):
    def __init__(self, car_dimensions, sample_time: float, initial_state: State):
        super().__init__(car_dimensions, sample_time, initial_state)
        self.history = History(sample_time=sample_time)
        self.history.store(initial_state, a=0., delta=0., xref_deviation=0.)

    def step(self, a: float, delta: float, xref_deviation: Optional[float] = None) -> State:
        new_state = super().step(a, delta)
        self.history.store(new_state, a=a, delta=delta, xref_deviation=xref_deviation)
        return new_state
