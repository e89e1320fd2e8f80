"""m17_cxx_demod_b200 -- B200-native many-channel M17 demodulator (hot path of mobilinkd/m17-cxx-demod).

The product is the CUDA library ``libm17b.so`` (C ABI in ``include/m17b.h``); this
package is a thin ctypes binding used by the tests and by ``bench.py``.  There is
no CPU fallback: without the built library or without a CUDA device every entry
point raises.
"""
from .demod import (Demodulator, FRAME_DTYPE, FRAME_TYPE_NAMES, PAYLOAD_LEN, FIR_FMA, INVERT, PROFILE, STREAMING,
                    M17BError, library_path, load_library, device_count)

__all__ = ["Demodulator", "FRAME_DTYPE", "FRAME_TYPE_NAMES", "PAYLOAD_LEN", "FIR_FMA", "INVERT", "PROFILE", "STREAMING",
           "M17BError", "library_path", "load_library", "device_count"]
