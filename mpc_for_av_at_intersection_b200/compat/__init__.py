"""Reference-compatible faces of the batched engine: modules shaped like the reference's `lib.mpc`, `lib.mpc_with_speed`, `lib.mpc_sensitivity`, `lib.mpc_jerk`,
`lib.simulation` and `lib.car_dimensions` so that main/scenarios/*.py run by changing only their imports
(see INTEGRATION.md).  `install()` aliases them under the reference's module names."""
import sys


def install(prefixes=("lib", "main.lib")):
    """Register the compat modules as `lib.mpc`, `lib.simulation`, `lib.car_dimensions` (and `main.lib.*`) in
    sys.modules, the way a maintainer of the reference would switch the scenario scripts over."""
    from . import car_dimensions, mpc, mpc_jerk, mpc_sensitivity, mpc_with_speed, simulation
    for pre in prefixes:
        sys.modules[f"{pre}.mpc"] = mpc
        sys.modules[f"{pre}.mpc_with_speed"] = mpc_with_speed
        sys.modules[f"{pre}.mpc_sensitivity"] = mpc_sensitivity
        sys.modules[f"{pre}.mpc_jerk"] = mpc_jerk
        sys.modules[f"{pre}.simulation"] = simulation
        sys.modules[f"{pre}.car_dimensions"] = car_dimensions
    return mpc
