"""B200-native batched MPC tracking loop behind the reference's `MPC` API."""
from .config import MpcConfig  # noqa: F401

__all__ = ["MpcConfig"]
