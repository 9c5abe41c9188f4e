"""B200-native batched MPC tracking loop behind the reference's `MPC` API."""
from .config import MpcConfig  # noqa: F401
from .courses import CourseSet, load_intersection_routes  # noqa: F401
from .session import MpcSession  # noqa: F401

__all__ = ["MpcConfig", "CourseSet", "MpcSession", "load_intersection_routes"]
