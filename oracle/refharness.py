"""TEST INFRASTRUCTURE -- import the UNMODIFIED reference (read-only, /root/reference).

Only usable in the build container (the GPU box has no /root/reference); it is
used by oracle/gen_golden.py to produce tests/golden/*.npz and by the CPU tests
that validate the numpy/C restatements against the reference itself.

Shims, none of which touches the reference tree:
  1. a stub `matplotlib` package (main/bicycle/main.py:1 imports pyplot at
     module scope and main/lib/simulation.py:7 imports bicycle.main);
  2. a replacement `lib.motion_primitive` whose dataclass uses default_factory
     (main/lib/motion_primitive.py:19 has a mutable default, rejected by py3.12);
  3. oracle/cvxpy_shim on sys.path as `cvxpy`;
  4. cwd = <scratch>/scenarios with <scratch>/config/mpc_config.json, because
     main/lib/mpc.py:13 opens '../config/mpc_config.json' relative to cwd.
"""
import importlib
import json
import os
import pickle
import sys
import tempfile
import types
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

REF_ROOT = Path(os.environ.get("MPC_REFERENCE_ROOT", "/root/reference"))
ORACLE_DIR = Path(__file__).resolve().parent


def available() -> bool:
    return (REF_ROOT / "main" / "lib" / "mpc.py").exists()


class _Anything(types.ModuleType):
    """Module whose every attribute is a no-op callable/namespace (matplotlib stub)."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        sub = _Anything(self.__name__ + "." + name)
        setattr(self, name, sub)
        return sub

    def __call__(self, *a, **k):
        return _Anything("call")

    def __iter__(self):
        return iter(())

    def __mro_entries__(self, bases):
        return (object,)


def _install_matplotlib_stub():
    if "matplotlib" in sys.modules and not isinstance(sys.modules["matplotlib"], _Anything):
        return
    root = _Anything("matplotlib")
    root.__path__ = []
    sys.modules["matplotlib"] = root
    for sub in ("pyplot", "patches", "colors", "lines", "animation", "ticker", "collections",
                "transforms", "cm", "axes", "figure"):
        m = _Anything("matplotlib." + sub)
        sys.modules["matplotlib." + sub] = m
        setattr(root, sub, m)


def _install_motion_primitive_module():
    mod = types.ModuleType("lib.motion_primitive")

    @dataclass
    class MotionPrimitive:
        name: str
        forward_speed: float
        steering_angle: float
        n_seconds: float
        total_length: float = 0.
        points: np.ndarray = field(default_factory=lambda: np.array([]))

    MotionPrimitive.__module__ = "lib.motion_primitive"
    mod.MotionPrimitive = MotionPrimitive

    def load_motion_primitives(version="prius"):
        d = REF_ROOT / "main" / "data" / ("motion_primitives_" + version)
        out = {}
        for fn in sorted(d.glob("*.pkl")):
            with open(fn, "rb") as f:
                mp = pickle.load(f)
            out[mp.name] = mp
        return out

    mod.load_motion_primitives = load_motion_primitives
    sys.modules["lib.motion_primitive"] = mod
    sys.modules["main.lib.motion_primitive"] = mod


_scratch = None


def setup(config_overrides=None):
    """Install shims, chdir into a scratch dir with the desired mpc_config.json.

    Returns the parsed config dict.  Call `load_mpc()` afterwards to (re)import lib.mpc.
    """
    global _scratch
    assert available(), "reference tree not present"
    sys.dont_write_bytecode = True
    _install_matplotlib_stub()
    for p in (str(ORACLE_DIR / "cvxpy_shim"), str(REF_ROOT / "main"), str(REF_ROOT)):
        if p not in sys.path:
            sys.path.insert(0, p)
    # `lib` must resolve to main/lib; make the package known, then patch in the shimmed submodule
    importlib.import_module("lib")
    importlib.import_module("main.lib")
    _install_motion_primitive_module()

    with open(REF_ROOT / "main" / "config" / "mpc_config.json") as f:
        cfg = json.load(f)
    cfg.update(config_overrides or {})
    _scratch = tempfile.mkdtemp(prefix="mpcref_")
    os.makedirs(os.path.join(_scratch, "config"))
    os.makedirs(os.path.join(_scratch, "scenarios"))
    with open(os.path.join(_scratch, "config", "mpc_config.json"), "w") as f:
        json.dump(cfg, f)
    with open(os.path.join(_scratch, "config", "mpc_config_sensitivity.json"), "w") as f:
        json.dump(cfg, f)
    os.chdir(os.path.join(_scratch, "scenarios"))
    return cfg


def load_mpc(module="lib.mpc"):
    """(Re)import the reference MPC module so it re-reads the scratch config."""
    if module in sys.modules:
        del sys.modules[module]
    return importlib.import_module(module)


def plan_route(start_pos: int, turn_indicator: int, version="bicycle_model"):
    """Run the reference's own planner (main/scenarios/mpc_intersection.py:31-60) for one route."""
    from envs.intersection import intersection
    from lib.car_dimensions import BicycleModelDimensions
    from lib.motion_primitive import load_motion_primitives
    from lib.motion_primitive_search_modified import MotionPrimitiveSearch
    car = BicycleModelDimensions(skip_back_circle_collision_checking=False)
    mps = load_motion_primitives(version=version)
    scenario = intersection(start_pos=start_pos, turn_indicator=turn_indicator)
    search = MotionPrimitiveSearch(scenario, car, mps, margin=car.radius)
    cost, path, traj = search.run(debug=False)
    return cost, path, traj
