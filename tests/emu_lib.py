"""ctypes access to tests/emu/libm17emu.so: the HOST build of the kernels' per-lane logic
(m17_cxx_demod_b200/csrc/m17b_lane.cuh compiled with g++).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        subprocess.run(["make", "-s"], cwd=os.path.join(HERE, "emu"), check=True)
        _lib = ctypes.CDLL(os.path.join(HERE, "emu", "libm17emu.so"))
        _lib.m17emu_demod.restype = ctypes.c_long
    return _lib


def demod(samples: np.ndarray, chunks, invert: bool = False, rec_dtype=None, streaming: bool = False, origin: int = 0,
          split_seed: int | None = None) -> list[np.ndarray]:
    """Run the emulated scan/decode lanes over [n_streams, n] int16 fed in submits of the given sizes.  split_seed: emulate
    M17B_STREAMING's two independent chains of rounds (receiving / acquiring streams) in a seeded arbitrary interleaving."""
    lib().m17emu_set_origin(ctypes.c_uint32(origin & 0xFFFFFFFF))
    from m17_cxx_demod_b200.demod import FRAME_DTYPE, split_by_stream
    samples = np.ascontiguousarray(samples, dtype=np.int16)
    ns, n = samples.shape
    cs = np.asarray(list(chunks), dtype=np.uint32)
    cap = ns * (2 * (n // 1600 + 2) + 8)
    out = np.zeros(cap, dtype=FRAME_DTYPE)
    k = lib().m17emu_demod(samples.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(ns), ctypes.c_size_t(n), ctypes.c_size_t(n),
                           cs.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(len(cs)), ctypes.c_int(int(invert) | (2 if streaming else 0) | ((4 | ((split_seed & 0xFFFF) << 8)) if split_seed is not None else 0)),
                           out.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(cap))
    if k < 0:
        raise RuntimeError(f"m17emu_demod failed: {k}")
    return split_by_stream(out[:k], ns)
