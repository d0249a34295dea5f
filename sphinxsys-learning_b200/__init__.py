"""sphb200: B200-native (sm_100a) implementation of SPHinXsys's particle-interaction hot path.

The product is libsphb200.so (hand-written CUDA behind the C ABI in include/sphb200.h).  This
package holds its sources (csrc/), the ctypes binding (capi), the host-side mirror of the reference's
interface used by tests and bench (host) and the reference's benchmark cases as data (cases).
"""
from . import capi, cases, host  # noqa: F401
from .host import SPHSystem  # noqa: F401
