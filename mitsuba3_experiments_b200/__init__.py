"""b2path: B200-native path tracer + PRB adjoint behind Mitsuba 3's integrator interface."""
from . import capi  # noqa: F401
from .scene import SceneDesc, T  # noqa: F401
from .scenes import cornell_box  # noqa: F401

__all__ = ["capi", "SceneDesc", "T", "cornell_box"]
