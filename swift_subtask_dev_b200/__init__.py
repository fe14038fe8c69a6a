"""B200-native SPH neighbour-loop backend (density + ghost + force) behind the
C ABI of include/swift_cuda.h.  See DESIGN.md."""
from . import ics  # noqa: F401
from .backend import HydroBackend, make_params, build_library, load_library, SwiftCudaError  # noqa: F401
