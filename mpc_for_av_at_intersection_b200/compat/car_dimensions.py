"""Vehicle geometry with the attribute names of the reference's CarDimensions hierarchy
(main/lib/car_dimensions.py:7-107).  Only `distance_back_to_front_wheel` enters the MPC (main/lib/mpc.py:154);
`radius` and `circle_centers` feed the collision check (main/lib/collision_avoidance.py:66-104)."""
from dataclasses import dataclass
from typing import Tuple

import numpy as np


@dataclass(frozen=True)
class CarDimensions:
    """Rear-axle-anchored rectangle approximated by two circles."""
    wheelbase: float
    width: float
    length: float
    skip_back_circle_collision_checking: bool = False

    @property
    def distance_back_to_front_wheel(self) -> float:
        return self.wheelbase

    @property
    def bounding_box_size(self) -> Tuple[float, float]:
        return self.width, self.length

    @property
    def center_point_offset(self) -> Tuple[float, float]:
        return self.wheelbase / 2, 0.0                      # car_dimensions.py:59-60

    @property
    def radius(self) -> float:
        return self.width / (2 ** .5)                       # car_dimensions.py:64-65

    @property
    def circle_centers(self) -> np.ndarray:
        along = self.length / 2 - self.width / 2            # car_dimensions.py:69-79
        cx, cy = self.center_point_offset
        rows = [[cx + along, cy]]
        if not self.skip_back_circle_collision_checking:
            rows.append([cx - along, cy])
        return np.array(rows)


def BicycleModelDimensions(skip_back_circle_collision_checking=False) -> CarDimensions:
    """L = 2.86, box 2.0 x (L + 0.64)  (car_dimensions.py:82-90)."""
    return CarDimensions(2.86, 2.0, 2.86 + 0.64, skip_back_circle_collision_checking)


def PriusDimensions(scaling_factor: float = 1., skip_back_circle_collision_checking=False) -> CarDimensions:
    """L = 4 s, box 2.04 s x 4.84 s  (car_dimensions.py:93-107)."""
    return CarDimensions(4 * scaling_factor, 2.04 * scaling_factor, 4.84 * scaling_factor,
                         skip_back_circle_collision_checking)
