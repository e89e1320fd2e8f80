"""ctypes binding of include/m17b.h (libm17b.so)."""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

FIR_FMA, INVERT, PROFILE, STREAMING, DIAG_RECORDS = 0x1, 0x2, 0x4, 0x8, 0x10
HOST, DEVICE = 0, 1

FRAME_DTYPE = np.dtype([
    ("stream", "<u4"), ("seq", "<u4"), ("sample_pos", "<u8"), ("viterbi_cost", "<i4"),
    ("type", "u1"), ("pad", "u1", (3,)), ("payload", "u1", (26,)), ("lsf", "u1", (30,)),
])
assert FRAME_DTYPE.itemsize == 80
DIAG_REC_DTYPE = np.dtype([
    ("stream", "<u4"), ("pad", "<u4"), ("dcd", "<i4"), ("evm", "<f4"), ("deviation", "<f4"), ("offset", "<f4"), ("locked", "<i4"),
    ("clock", "<f4"), ("sample_index", "<i4"), ("sync_index", "<i4"), ("clock_index", "<i4"), ("viterbi_cost", "<i4"),
    ("samples", "<u8"),
])
assert DIAG_REC_DTYPE.itemsize == 56
FRAME_TYPE_NAMES = ["LSF", "LICH", "STREAM", "BASIC_PACKET", "FULL_PACKET", "BERT"]
PAYLOAD_LEN = {0: 0, 1: 6, 2: 18, 3: 26, 4: 26, 5: 25}


class M17BError(RuntimeError):
    pass


class _Config(ctypes.Structure):
    _fields_ = [("n_streams", ctypes.c_uint32), ("device", ctypes.c_int32), ("max_chunk", ctypes.c_uint32),
                ("flags", ctypes.c_uint32), ("frame_capacity", ctypes.c_uint32), ("reserved", ctypes.c_uint32 * 3)]


class Diag(ctypes.Structure):
    _fields_ = [("dcd", ctypes.c_int32), ("evm", ctypes.c_float), ("deviation", ctypes.c_float),
                ("offset", ctypes.c_float), ("locked", ctypes.c_int32), ("clock", ctypes.c_float),
                ("sample_index", ctypes.c_int32), ("sync_index", ctypes.c_int32), ("clock_index", ctypes.c_int32),
                ("viterbi_cost", ctypes.c_int32), ("samples", ctypes.c_uint64)]


class Counters(ctypes.Structure):
    _fields_ = [("samples", ctypes.c_uint64), ("frames", ctypes.c_uint64 * 6), ("lsf_crc_fail", ctypes.c_uint64),
                ("golay_fail", ctypes.c_uint64), ("faults", ctypes.c_uint64), ("dropped", ctypes.c_uint64),
                ("packets_ok", ctypes.c_uint64), ("packets_crc_fail", ctypes.c_uint64), ("packet_seq_errors", ctypes.c_uint64),
                ("packets_dropped", ctypes.c_uint64)]


class Ber(ctypes.Structure):
    _fields_ = [("synced", ctypes.c_int32), ("bits", ctypes.c_uint32), ("errors", ctypes.c_uint32), ("packet_bytes", ctypes.c_uint32)]


PACKET_MAX = 860
PACKET_DTYPE = np.dtype([("stream", "<u4"), ("seq", "<u4"), ("sample_pos", "<u8"), ("length", "<u2"), ("fcs_ok", "u1"), ("pad", "u1"),
                         ("data", "u1", (PACKET_MAX,))])
assert PACKET_DTYPE.itemsize == 880


class Profile(ctypes.Structure):
    _fields_ = [("fir_ms", ctypes.c_double), ("fir_launches", ctypes.c_uint64), ("fir_samples", ctypes.c_uint64),
                ("scan_ms", ctypes.c_double), ("scan_launches", ctypes.c_uint64), ("scan_samples", ctypes.c_uint64),
                ("decode_ms", ctypes.c_double), ("decode_launches", ctypes.c_uint64), ("decode_frames", ctypes.c_uint64),
                ("aux_ms", ctypes.c_double), ("aux_launches", ctypes.c_uint64),
                ("h2d_ms", ctypes.c_double), ("h2d_bytes", ctypes.c_uint64),
                ("d2h_ms", ctypes.c_double), ("d2h_bytes", ctypes.c_uint64),
                ("dcd_ms", ctypes.c_double), ("dcd_launches", ctypes.c_uint64),
                ("fixup_ms", ctypes.c_double), ("fixup_launches", ctypes.c_uint64),
                ("search_requests", ctypes.c_uint64), ("search_samples", ctypes.c_uint64), ("scan1_ms", ctypes.c_double)]


_lib = None


def library_path() -> str:
    return os.path.join(_HERE, "libm17b.so")


def load_library() -> ctypes.CDLL:
    """Load libm17b.so; raises (loudly) when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise M17BError(f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU fallback)")
        lib = ctypes.CDLL(path)
        lib.m17b_strerror.restype = ctypes.c_char_p
        lib.m17b_last_cuda_error.restype = ctypes.c_char_p
        lib.m17b_last_cuda_error.argtypes = [ctypes.c_void_p]
        lib.m17b_version.restype = ctypes.c_uint32
        for name in ("m17b_create", "m17b_submit", "m17b_sync", "m17b_drain", "m17b_pending_frames", "m17b_diag_get",
                     "m17b_counters_get", "m17b_get_profile", "m17b_debug_filter", "m17b_debug_correlate", "m17b_diag_drain",
                     "m17b_measure_fp32_peak",
                     "m17b_host_alloc", "m17b_device_alloc", "m17b_memcpy_h2d", "m17b_device_count",
                     "m17b_event_record", "m17b_event_elapsed", "m17b_profile_enable", "m17b_packet_drain", "m17b_ber_get", "m17b_filter_noalloc"):
            getattr(lib, name).restype = ctypes.c_int
        lib.m17b_submit.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int]
        lib.m17b_sync.argtypes = [ctypes.c_void_p]
        lib.m17b_drain.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
        lib.m17b_pending_frames.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_size_t)]
        lib.m17b_ticket.restype = ctypes.c_ulonglong
        lib.m17b_ticket.argtypes = [ctypes.c_void_p]
        lib.m17b_drain_upto.restype = ctypes.c_int
        lib.m17b_drain_upto.argtypes = [ctypes.c_void_p, ctypes.c_ulonglong, ctypes.c_void_p, ctypes.c_size_t,
                                        ctypes.POINTER(ctypes.c_size_t)]
        lib.m17b_destroy.argtypes = [ctypes.c_void_p]
        lib.m17b_destroy.restype = None
        lib.m17b_host_alloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t]
        lib.m17b_host_free.argtypes = [ctypes.c_void_p]
        lib.m17b_host_free.restype = None
        lib.m17b_device_alloc.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t]
        lib.m17b_device_free.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.m17b_device_free.restype = None
        lib.m17b_memcpy_h2d.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
        lib.m17b_diag_get.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.POINTER(Diag)]
        lib.m17b_counters_get.argtypes = [ctypes.c_void_p, ctypes.POINTER(Counters)]
        lib.m17b_get_profile.argtypes = [ctypes.c_void_p, ctypes.POINTER(Profile), ctypes.c_int]
        lib.m17b_debug_filter.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t,
                                          ctypes.c_void_p, ctypes.c_void_p]
        lib.m17b_debug_correlate.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t,
                                             ctypes.c_void_p, ctypes.c_void_p]
        lib.m17b_diag_drain.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
        lib.m17b_filter_noalloc.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_size_t, ctypes.c_size_t,
                                            ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p]
        lib.m17b_profile_enable.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.m17b_packet_drain.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
        lib.m17b_ber_get.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.POINTER(Ber)]
        lib.m17b_event_record.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.m17b_event_elapsed.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
        lib.m17b_measure_fp32_peak.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
        _lib = lib
    return _lib


def device_count() -> int:
    return int(load_library().m17b_device_count())


class PinnedBuffer:
    """int16 host buffer in page-locked memory (cudaHostAlloc through the C ABI)."""

    def __init__(self, shape):
        self._lib = load_library()
        n = int(np.prod(shape))
        self._ptr = ctypes.c_void_p()
        rc = self._lib.m17b_host_alloc(ctypes.byref(self._ptr), ctypes.c_size_t(n * 2))
        if rc:
            raise M17BError(f"m17b_host_alloc({n * 2}) failed: {rc}")
        buf = (ctypes.c_int16 * n).from_address(self._ptr.value)
        self.array = np.frombuffer(buf, dtype=np.int16).reshape(shape)

    def free(self):
        if self._ptr:
            self.array = None
            self._lib.m17b_host_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Demodulator:
    """N resident M17 streams on one GPU (m17b_ctx)."""

    def __init__(self, n_streams: int, max_chunk: int, flags: int = 0, device: int = 0, frame_capacity: int = 0,
                 debug_origin: int = 0):
        self._lib = load_library()
        cfg = _Config(n_streams=n_streams, device=device, max_chunk=max_chunk, flags=flags,
                      frame_capacity=frame_capacity)
        cfg.reserved[0] = debug_origin & 0xFFFFFFFF      # test hook: ring position of the first sample
        self._ctx = ctypes.c_void_p()
        rc = self._lib.m17b_create(ctypes.byref(cfg), ctypes.byref(self._ctx))
        if rc:
            self._ctx = None
            raise M17BError(f"m17b_create failed: {self._lib.m17b_strerror(rc).decode()} ({rc})")
        self.n_streams = n_streams
        self.max_chunk = max_chunk
        self.flags = flags
        self._keep = []

    def _check(self, rc: int, what: str):
        if rc:
            msg = self._lib.m17b_strerror(rc).decode()
            detail = self._lib.m17b_last_cuda_error(self._ctx).decode() if self._ctx else ""
            raise M17BError(f"{what}: {msg} ({rc}) {detail}")

    def close(self):
        if self._ctx:
            self._lib.m17b_destroy(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def submit(self, samples: np.ndarray):
        """samples: int16 [n_streams, n] (last axis contiguous); asynchronous."""
        assert samples.dtype == np.int16 and samples.ndim == 2 and samples.shape[0] == self.n_streams
        assert samples.strides[1] == 2
        self._keep.append(samples)
        self._check(self._lib.m17b_submit(self._ctx, samples.ctypes.data_as(ctypes.c_void_p), samples.shape[1],
                                          samples.strides[0] // 2, HOST), "m17b_submit")

    def submit_device(self, dev_ptr: int, n: int, stride: int):
        self._check(self._lib.m17b_submit(self._ctx, ctypes.c_void_p(dev_ptr), n, stride, DEVICE), "m17b_submit")

    def sync(self):
        self._check(self._lib.m17b_sync(self._ctx), "m17b_sync")
        self._keep.clear()

    def _drain(self, ticket) -> np.ndarray:
        parts = []
        cap = max(4096, getattr(self, "_drain_cap", 4096))
        while True:
            out = np.empty(cap, dtype=FRAME_DTYPE)
            got = ctypes.c_size_t()
            if ticket is None:
                rc = self._lib.m17b_drain(self._ctx, out.ctypes.data_as(ctypes.c_void_p), cap, ctypes.byref(got))
            else:
                rc = self._lib.m17b_drain_upto(self._ctx, ctypes.c_ulonglong(ticket), out.ctypes.data_as(ctypes.c_void_p), cap,
                                               ctypes.byref(got))
            self._check(rc, "m17b_drain")
            parts.append(out[:got.value])
            if got.value < cap:
                break
            cap *= 4
        self._drain_cap = min(max(4096, 2 * sum(len(p) for p in parts)), 1 << 24)
        return parts[0] if len(parts) == 1 else np.concatenate(parts)

    def drain(self) -> np.ndarray:
        """All frame records decoded so far (chronological within each stream).  Waits for the device."""
        r = self._drain(None)
        self._keep.clear()
        return r

    def ticket(self) -> int:
        """Ticket of the most recent submit (see drain_upto)."""
        return int(self._lib.m17b_ticket(self._ctx))

    def drain_upto(self, ticket: int) -> np.ndarray:
        """Records appended until the submit with `ticket` finished; later submits keep running."""
        r = self._drain(ticket)
        if len(self._keep) > 8:
            del self._keep[:-8]
        return r

    def demodulate(self, samples: np.ndarray, chunk: int | None = None) -> list[np.ndarray]:
        """Feed [n_streams, n] int16 in chunks; return per-stream record arrays."""
        n = samples.shape[1]
        chunk = min(chunk or self.max_chunk, self.max_chunk)
        parts = []
        for o in range(0, n, chunk):
            self.submit(samples[:, o:o + chunk])
            parts.append(self.drain())
        rec = np.concatenate(parts) if parts else np.zeros(0, FRAME_DTYPE)
        return split_by_stream(rec, self.n_streams)

    def diag(self, stream: int) -> Diag:
        d = Diag()
        self._check(self._lib.m17b_diag_get(self._ctx, stream, ctypes.byref(d)), "m17b_diag_get")
        return d

    def counters(self) -> Counters:
        c = Counters()
        self._check(self._lib.m17b_counters_get(self._ctx, ctypes.byref(c)), "m17b_counters_get")
        return c

    def profile(self, reset: bool = False) -> Profile:
        p = Profile()
        self._check(self._lib.m17b_get_profile(self._ctx, ctypes.byref(p), int(reset)), "m17b_get_profile")
        return p

    def packet_drain(self) -> np.ndarray:
        """All packet records (one per final packet segment decoded) since the last call, m17b_packet_drain."""
        parts = []
        while True:
            buf = np.zeros(256, PACKET_DTYPE)
            n = ctypes.c_size_t()
            self._check(self._lib.m17b_packet_drain(self._ctx, buf.ctypes.data_as(ctypes.c_void_p), 256, ctypes.byref(n)), "m17b_packet_drain")
            parts.append(buf[:n.value])
            if n.value < 256:
                break
        return np.concatenate(parts)

    def ber(self, stream: int) -> Ber:
        b = Ber()
        self._check(self._lib.m17b_ber_get(self._ctx, stream, ctypes.byref(b)), "m17b_ber_get")
        return b

    def profile_enable(self, on: bool):
        self._check(self._lib.m17b_profile_enable(self._ctx, int(on)), "m17b_profile_enable")

    def debug_filter(self, samples: np.ndarray):
        """Matched filter only: returns (filtered, scaled) float32 [n_streams, n]."""
        samples = np.ascontiguousarray(samples, dtype=np.int16)
        assert samples.shape[0] == self.n_streams
        n = samples.shape[1]
        y = np.zeros((self.n_streams, n), dtype=np.float32)
        x = np.zeros((self.n_streams, n), dtype=np.float32)
        self._check(self._lib.m17b_debug_filter(self._ctx, samples.ctypes.data_as(ctypes.c_void_p), n, n,
                                                y.ctypes.data_as(ctypes.c_void_p), x.ctypes.data_as(ctypes.c_void_p)),
                    "m17b_debug_filter")
        return y, x

    def debug_correlate(self, samples: np.ndarray):
        """Matched filter + correlator clocked on every sample of fresh streams: (filtered [n_streams, n],
        corr [n_streams, n, 5] = limit_, preamble, LSF/stream, packet/BERT, EOT correlations)."""
        samples = np.ascontiguousarray(samples, dtype=np.int16)
        assert samples.shape[0] == self.n_streams
        n = samples.shape[1]
        y = np.zeros((self.n_streams, n), dtype=np.float32)
        c = np.zeros((self.n_streams, n, 5), dtype=np.float32)
        self._check(self._lib.m17b_debug_correlate(self._ctx, samples.ctypes.data_as(ctypes.c_void_p), n, n,
                                                   y.ctypes.data_as(ctypes.c_void_p), c.ctypes.data_as(ctypes.c_void_p)),
                    "m17b_debug_correlate")
        return y, c

    def diag_drain(self) -> np.ndarray:
        """All diagnostic records so far (contexts created with DIAG_RECORDS), chronological within each stream."""
        parts = []
        cap = 4096
        while True:
            out = np.empty(cap, dtype=DIAG_REC_DTYPE)
            got = ctypes.c_size_t()
            self._check(self._lib.m17b_diag_drain(self._ctx, out.ctypes.data_as(ctypes.c_void_p), cap, ctypes.byref(got)),
                        "m17b_diag_drain")
            parts.append(out[:got.value])
            if got.value < cap:
                break
        return np.concatenate(parts)

    def measure_fp32_peak(self, variant: int = 0) -> float:
        v = ctypes.c_double()
        self._check(self._lib.m17b_measure_fp32_peak(self._ctx, variant, ctypes.byref(v)), "m17b_measure_fp32_peak")
        return v.value

    def event_record(self, slot: int):
        self._check(self._lib.m17b_event_record(self._ctx, slot), "m17b_event_record")

    def event_elapsed_ms(self, a: int, b: int) -> float:
        v = ctypes.c_double()
        self._check(self._lib.m17b_event_elapsed(self._ctx, a, b, ctypes.byref(v)), "m17b_event_elapsed")
        return v.value

    def device_alloc(self, nbytes: int) -> int:
        p = ctypes.c_void_p()
        self._check(self._lib.m17b_device_alloc(self._ctx, ctypes.byref(p), nbytes), "m17b_device_alloc")
        return p.value

    def device_free(self, ptr: int):
        self._lib.m17b_device_free(self._ctx, ctypes.c_void_p(ptr))

    def memcpy_h2d(self, dev_ptr: int, src: np.ndarray):
        src = np.ascontiguousarray(src)
        self._check(self._lib.m17b_memcpy_h2d(self._ctx, ctypes.c_void_p(dev_ptr), src.ctypes.data_as(ctypes.c_void_p),
                                              src.nbytes), "m17b_memcpy_h2d")


def split_by_stream(rec: np.ndarray, n_streams: int) -> list[np.ndarray]:
    """Group records by stream (stable: keeps each stream's chronological order)."""
    rec = rec[np.lexsort((rec["seq"], rec["stream"]))]
    idx = np.searchsorted(rec["stream"], np.arange(n_streams + 1))
    return [rec[idx[s]:idx[s + 1]] for s in range(n_streams)]
