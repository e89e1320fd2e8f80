"""ctypes access to the CPU oracles (oracle/).  TEST INFRASTRUCTURE ONLY.

`ref()`  -> oracle/_ref/libm17ref.so : the unmodified reference headers.
`port()` -> oracle/libm17port.so     : the plain-C restatement.
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

RECORD_DTYPE = np.dtype([
    ("stream", "<u4"), ("seq", "<u4"), ("sample_pos", "<u8"), ("viterbi_cost", "<i4"),
    ("type", "u1"), ("pad", "u1", (3,)), ("payload", "u1", (26,)), ("lsf", "u1", (30,)),
])
assert RECORD_DTYPE.itemsize == 80

TRACE_DTYPE = np.dtype([
    ("filtered", "<f4"), ("limit", "<f4"), ("dcd_level", "<f4"), ("dev_offset", "<f4"), ("dev_idev", "<f4"),
    ("clock_est", "<f4"), ("sample_est", "<f4"), ("evm", "<f4"),
    ("sync_count", "<i4"), ("missing_sync_count", "<i4"), ("viterbi_cost", "<i4"), ("count", "<u4"),
    ("demod_state", "u1"), ("sample_index", "u1"), ("sync_sample_index", "u1"), ("clock_sample_index", "u1"),
    ("dcd", "u1"), ("dcd_triggered", "u1"), ("sync_word_type", "u1"), ("decoder_state", "u1"),
    ("corr_index", "u1"), ("need_clock_reset", "u1"), ("need_clock_update", "u1"), ("polarity", "u1"),
])
assert TRACE_DTYPE.itemsize == 60

PAYLOAD_LEN = {0: 0, 1: 6, 2: 18, 3: 26, 4: 26, 5: 25}
TYPE_NAMES = ["LSF", "LICH", "STREAM", "BASIC_PACKET", "FULL_PACKET", "BERT"]

_libs: dict[str, ctypes.CDLL] = {}


def _load(path: str) -> ctypes.CDLL:
    if path not in _libs:
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        _libs[path] = ctypes.CDLL(path)
    return _libs[path]


def ref_path() -> str:
    return os.path.join(ROOT, "oracle", "_ref", "libm17ref.so")


def ref_binary() -> str:
    return os.path.join(ROOT, "oracle", "_ref", "m17-demod-ref")


def have_ref() -> bool:
    return os.path.exists(ref_path())


def ref() -> ctypes.CDLL:
    return _load(ref_path())


def port() -> ctypes.CDLL:
    return _load(os.path.join(ROOT, "oracle", "libm17port.so"))


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def demod_batch(which: str, samples: np.ndarray, flags: int = 0, trace: bool = False, jobs: int | None = None,
                rec_cap: int | None = None):
    """Run an oracle over [n_streams, n_samples] int16.  Returns (list of record arrays, trace or None)."""
    samples = np.ascontiguousarray(samples, dtype=np.int16)
    if samples.ndim == 1:
        samples = samples[None, :]
    n_streams, n_samples = samples.shape
    if rec_cap is None:
        rec_cap = n_samples // 960 + 64
    rec = np.zeros((n_streams, rec_cap), dtype=RECORD_DTYPE)
    counts = np.zeros(n_streams, dtype=np.uint64)
    tr = np.zeros((n_streams, n_samples), dtype=TRACE_DTYPE) if trace else None
    if which == "ref":
        fn = ref().m17ref_demod_batch
    elif which == "port":
        fn = port().m17port_demod_batch
    else:
        raise ValueError(which)
    fn.restype = ctypes.c_int
    if jobs is None:
        jobs = os.cpu_count() or 1
    rc = fn(_p(samples), ctypes.c_size_t(n_streams), ctypes.c_size_t(n_samples), ctypes.c_size_t(n_samples),
            ctypes.c_uint(flags), _p(rec), ctypes.c_size_t(rec_cap), _p(counts),
            _p(tr) if trace else None, ctypes.c_int(jobs))
    if rc:
        raise RuntimeError(f"{which} oracle batch failed rc={rc}")
    out = []
    for s in range(n_streams):
        n = int(counts[s])
        assert n <= rec_cap, "record capacity exceeded"
        out.append(rec[s, :n].copy())
    return out, tr


def records_equal(a: np.ndarray, b: np.ndarray, check_pos: bool = True) -> bool:
    """Bit-exact comparison of two record arrays (all fields incl. stale union bytes)."""
    if len(a) != len(b):
        return False
    fields = ["seq", "viterbi_cost", "type", "payload", "lsf"] + (["sample_pos"] if check_pos else [])
    return all(np.array_equal(a[f], b[f]) for f in fields)


def describe_mismatch(a: np.ndarray, b: np.ndarray) -> str:
    n = min(len(a), len(b))
    for i in range(n):
        for f in ("sample_pos", "type", "viterbi_cost", "payload", "lsf"):
            if not np.array_equal(a[i][f], b[i][f]):
                return f"record {i} field {f}: {a[i][f]} vs {b[i][f]} (len {len(a)} vs {len(b)})"
    return f"lengths {len(a)} vs {len(b)}"
