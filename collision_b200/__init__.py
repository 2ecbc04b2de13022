"""collision_b200 -- B200-native LBVH sphere-collision path behind the Python
API of kwohlfahrt/collision (`import collision_b200 as collision`).

Submodules mirror the reference package one to one: collision, radix, scan,
reduce, bounds, summer, index, offset, misc.  All compute is hand-written
sm_100a CUDA in libcollision_b200.so (include/collision_b200.h); there is no
CPU or PyTorch fallback -- a missing library raises on first use.
"""
from . import misc  # noqa: F401

__version__ = "0.1.0"
__all__ = ["collision", "radix", "scan", "reduce", "bounds", "summer", "index", "offset", "misc"]
