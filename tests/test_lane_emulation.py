"""The kernels' per-lane logic (m17b_lane.cuh: block/event-pass scan, frame decode) compiled for the
HOST and run against the CPU oracle.  Checks the control flow the GPU executes -- dense pass running
ahead of the event pass, resume after a pending frame, ragged submits, carrier-detect gaps -- without
a GPU.  The GPU parity tests (test_gpu_parity.py) remain the parity tests proper."""
import numpy as np
import pytest

import emu_lib
import oracle_lib as ol
from m17_cxx_demod_b200 import gen


def random_params(seed, mode, n):
    rng = np.random.default_rng(seed)
    return [gen.default_params(seed=int(rng.integers(1, 1 << 40)), mode=mode, snr_db=float(rng.uniform(-2, 30)),
                               skew_ppm=float(rng.uniform(-500, 500)), dc_offset=float(rng.uniform(-400, 400)), repeat=1,
                               gap=int(rng.integers(2000, 9000)), lead_in=int(rng.integers(0, 9600)),
                               n_frames=int(rng.integers(0, 30)), packet_encapsulated=int(rng.integers(0, 2)))
            for _ in range(n)]


def check(want, got, label):
    bad = [f"{label} stream {s}: " + ol.describe_mismatch(a, b) for s, (a, b) in enumerate(zip(want, got))
           if not ol.records_equal(a, b)]
    assert not bad, f"{len(bad)} of {len(want)} streams differ; first: {bad[:3]}"


@pytest.mark.parametrize("mode", [gen.STREAM, gen.PACKET, gen.BERT, gen.MIXED])
def test_emulated_lanes_match_oracle(mode):
    x = gen.generate(random_params(120 + mode, mode, 24), 96000)
    want, _ = ol.demod_batch("port", x)
    got = emu_lib.demod(x, [24000])
    check(want, got, f"mode {mode}")
    assert sum(len(r) for r in want) > 50


def test_emulated_lanes_ragged_submits():
    x = gen.generate(random_params(7, gen.MIXED, 12), 48011)
    want, _ = ol.demod_batch("port", x)
    got = emu_lib.demod(x, [1, 7, 8, 13, 1919, 1920, 1921, 4800, 5003])
    check(want, got, "ragged")


@pytest.mark.parametrize("mode", [gen.STREAM, gen.MIXED])
def test_emulated_lanes_streaming_deferral(mode):
    """M17B_STREAMING: lanes leave partial frames for the next submit; records must not change."""
    x = gen.generate(random_params(300 + mode, mode, 16), 96000)
    want, _ = ol.demod_batch("port", x)
    for chunks in ([9600], [4800], [1920, 3000, 7777]):
        got = emu_lib.demod(x, chunks, streaming=True)
        check(want, got, f"streaming mode {mode} chunks {chunks}")


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_emulated_lanes_two_independent_chains(seed):
    """M17B_STREAMING as m17b_submit enqueues it: the receiving and the acquiring streams have a chain of rounds each,
    every round classifies for itself, a stream changes chains through its flags word (an acquiring stream stops at the
    start of a frame and after its walk budget; a receiving one that loses sync or needs the fix-up kernels becomes
    acquiring), and either chain may be up to two submits behind.  Whatever the interleaving, the records must not change."""
    x = gen.generate(random_params(400 + seed, gen.MIXED, 16), 96000)
    want, _ = ol.demod_batch("port", x)
    for chunks in ([9600], [4800, 2500], [1920, 3000, 7777]):
        got = emu_lib.demod(x, chunks, streaming=True, split_seed=seed)
        check(want, got, f"split chains seed {seed} chunks {chunks}")
    assert sum(len(r) for r in want) > 100


def test_emulated_lanes_across_the_32_bit_position_wrap():
    """Ring positions are uint32: a stream that starts 30,016 samples before 2^32 decodes across the wrap unchanged."""
    x = gen.generate(random_params(51, gen.MIXED, 12), 96000)
    want, _ = ol.demod_batch("port", x)
    for streaming, chunks in ((False, [4801]), (True, [9600])):
        got = emu_lib.demod(x, chunks, streaming=streaming, origin=2**32 - 30016)
        check(want, got, f"wrap streaming={streaming}")


def test_emulated_lanes_idle_and_zero_inputs():
    rng = np.random.default_rng(3)
    x = np.zeros((4, 30000), np.int16)
    x[1] = rng.normal(0, 200, 30000).astype(np.int16)
    x[2] = rng.normal(0, 8000, 30000).astype(np.int16)
    x[3, 10000:] = rng.normal(0, 50, 20000).astype(np.int16)
    want, _ = ol.demod_batch("port", x)
    got = emu_lib.demod(x, [30000])
    check(want, got, "idle")


def test_soft_slicer_lookup_matches_oracle():
    """llr4 (four look-ups around an arithmetic guess) against the oracle's binary search (Util.h:128-145):
    a dense sweep of [-3.2, 3.2], every threshold with its neighbouring floats, specials."""
    import ctypes
    port = ol.port()
    thr = np.zeros(43, np.float32); pairs = np.zeros(86, np.int8)
    port.m17port_llr_table(thr.ctypes.data_as(ctypes.c_void_p), pairs.ctypes.data_as(ctypes.c_void_p))
    near = np.concatenate([np.nextafter(thr, np.float32(-9)), thr, np.nextafter(thr, np.float32(9))])
    for _ in range(3):
        near = np.concatenate([near, np.nextafter(near, np.float32(-9)), np.nextafter(near, np.float32(9))])
    x = np.concatenate([np.linspace(-3.2, 3.2, 400001).astype(np.float32), near.astype(np.float32),
                        np.array([0.0, -0.0, 3.0, -3.0, 1e9, -1e9, np.inf, -np.inf], np.float32)])
    got = np.zeros(len(x), np.uint16)
    emu_lib.lib().m17emu_llr4(x.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(len(x)), got.ctypes.data_as(ctypes.c_void_p))
    want = np.zeros((len(x), 2), np.int8)
    port.m17port_llr.argtypes = [ctypes.c_float, ctypes.c_void_p]
    for i, s in enumerate(x):
        port.m17port_llr(float(s), want[i].ctypes.data_as(ctypes.c_void_p))
    w16 = want.view(np.uint8).astype(np.uint16)
    assert np.array_equal(got, w16[:, 0] | (w16[:, 1] << 8))
