"""ctypes binding of the synthetic M17 baseband generator (include/m17gen.h).

The generator restates the reference's transmit path (apps/m17-mod.cpp:204-564)
with random payloads and seeded impairments; it is CPU tooling for tests and
bench.py and is not on the demodulation path.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

STREAM, PACKET, BERT, MIXED = 0, 1, 2, 3


class GenParams(ctypes.Structure):
    _fields_ = [
        ("seed", ctypes.c_uint64), ("mode", ctypes.c_int32), ("n_frames", ctypes.c_int32),
        ("lead_in", ctypes.c_int32), ("tail", ctypes.c_int32), ("packet_encapsulated", ctypes.c_int32),
        ("invert", ctypes.c_int32), ("repeat", ctypes.c_int32), ("gap", ctypes.c_int32),
        ("snr_db", ctypes.c_float), ("skew_ppm", ctypes.c_float), ("dc_offset", ctypes.c_float),
        ("gain", ctypes.c_float), ("phase", ctypes.c_float), ("idle_noise_rms", ctypes.c_float),
    ]


class GenTruth(ctypes.Structure):
    _fields_ = [
        ("n_transmissions", ctypes.c_uint32), ("n_stream_frames", ctypes.c_uint32),
        ("n_packet_frames", ctypes.c_uint32), ("n_bert_frames", ctypes.c_uint32),
        ("first_lsf", ctypes.c_uint8 * 30), ("pad", ctypes.c_uint8 * 2),
    ]


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "libm17gen.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        _lib = ctypes.CDLL(path)
        _lib.m17gen_stream.restype = ctypes.c_int
        _lib.m17gen_batch.restype = ctypes.c_int
        _lib.m17gen_crc16.restype = ctypes.c_uint16
        _lib.m17gen_golay24_encode.restype = ctypes.c_uint32
    return _lib


def default_params(**kw) -> GenParams:
    p = GenParams()
    lib().m17gen_default_params(ctypes.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def generate(params: list[GenParams], n_samples: int, n_threads: int | None = None,
             out: np.ndarray | None = None) -> np.ndarray:
    """Generate len(params) streams -> int16 array [n_streams, n_samples] (stream-major)."""
    n = len(params)
    if out is None:
        out = np.empty((n, n_samples), dtype=np.int16)
    assert out.shape == (n, n_samples) and out.dtype == np.int16 and out.flags.c_contiguous
    arr = (GenParams * n)(*params)
    if n_threads is None:
        n_threads = min(os.cpu_count() or 1, max(1, n))
    rc = lib().m17gen_batch(arr, ctypes.c_size_t(n), out.ctypes.data_as(ctypes.c_void_p),
                            ctypes.c_size_t(n_samples), ctypes.c_size_t(n_samples), None, ctypes.c_int(n_threads))
    if rc:
        raise RuntimeError(f"m17gen_batch failed: {rc}")
    return out


_cuda_lib = None


def cuda_lib() -> ctypes.CDLL:
    """libm17gen_cuda.so: the device-side modulator (csrc/m17gen_dev.cu)."""
    global _cuda_lib
    if _cuda_lib is None:
        path = os.path.join(_HERE, "libm17gen_cuda.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        _cuda_lib = ctypes.CDLL(path)
        _cuda_lib.m17gen_batch_device.restype = ctypes.c_int
    return _cuda_lib


def generate_device(params: list[GenParams], n_samples: int, out_ptr: int, stride: int | None = None, device: int = 0) -> None:
    """Build len(params) streams of n_samples int16 ON THE GPU into device memory at out_ptr (stream i at
    out_ptr + 2 * i * stride): the same samples `generate` gives, bit for bit."""
    n = len(params)
    arr = (GenParams * n)(*params)
    lib().m17gen_snr_linear.restype = ctypes.c_double
    lib().m17gen_snr_linear.argtypes = [ctypes.c_float]
    snr = np.array([lib().m17gen_snr_linear(p.snr_db) for p in params], np.float64)
    rc = cuda_lib().m17gen_batch_device(arr, snr.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(n), ctypes.c_void_p(out_ptr),
                                        ctypes.c_size_t(n_samples), ctypes.c_size_t(stride or n_samples), ctypes.c_int(device))
    if rc:
        raise RuntimeError(f"m17gen_batch_device failed: {rc}")


@dataclass
class Workload:
    """A named synthetic corpus (SURVEY.md section 8d / BASELINE.json configs)."""
    name: str
    params: list
    n_samples: int


def stream_seed(config: int, stream: int) -> int:
    return 0x4D3137 ^ (config << 32) ^ stream


def workload_c1(seconds: float = 60.0) -> Workload:
    """configs[0]: one clean 60 s stream-mode transmission (SNR 40 dB so the DCD never sees exact zeros)."""
    p = default_params(seed=stream_seed(1, 0), mode=STREAM, snr_db=40.0)
    return Workload("c1_single_clean", [p], int(seconds * 48000))


def ebn0_to_snr_db(ebn0_db: float) -> float:
    # Eb/N0 = SNR * sps / 4 for real 4-FSK baseband at 10 samples/symbol, 2 bits/symbol.
    return ebn0_db - 3.9794


def workload_c2(n_streams: int = 1024, seconds: float = 10.0) -> Workload:
    """configs[1]: stream-mode channels, AWGN at Eb/N0 = 3..12 dB swept over streams."""
    ps = []
    for s in range(n_streams):
        ebn0 = 3.0 + 9.0 * (s % 10) / 9.0
        ps.append(default_params(seed=stream_seed(2, s), mode=STREAM, snr_db=ebn0_to_snr_db(ebn0)))
    return Workload("c2_awgn_ebn0_3_12", ps, int(seconds * 48000))


def workload_c3(n_unique: int = 4096, seconds: float = 2.0) -> Workload:
    """configs[2]: +-500 ppm clock skew and +-0.1 frequency offset (DC), SNR 20 dB."""
    rng = np.random.default_rng(0xC3)
    ps = []
    for s in range(n_unique):
        skew = float(rng.uniform(-500.0, 500.0))
        dc = float(rng.uniform(-0.1, 0.1)) * 41067.0 / 10.0069
        ps.append(default_params(seed=stream_seed(3, s), mode=STREAM, snr_db=20.0, skew_ppm=skew,
                                 dc_offset=dc, lead_in=4800 + int(rng.integers(0, 1920))))
    return Workload("c3_skew_offset", ps, int(seconds * 48000))


def workload_c4(n_unique: int = 4096, seconds: float = 2.0) -> Workload:
    """configs[3]: mixed LSF + stream + packet + BERT traffic, SNR 15 dB."""
    rng = np.random.default_rng(0xC4)
    ps = []
    for s in range(n_unique):
        ps.append(default_params(seed=stream_seed(4, s), mode=MIXED, snr_db=15.0, repeat=1,
                                 gap=4800 + int(rng.integers(0, 4800)),
                                 packet_encapsulated=int(rng.integers(0, 2)),
                                 lead_in=2400 + int(rng.integers(0, 4800))))
    return Workload("c4_mixed", ps, int(seconds * 48000))
