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


def test_emulated_lanes_idle_and_zero_inputs():
    rng = np.random.default_rng(3)
    x = np.zeros((4, 30000), np.int16)
    x[1] = rng.normal(0, 200, 30000).astype(np.int16)
    x[2] = rng.normal(0, 8000, 30000).astype(np.int16)
    x[3, 10000:] = rng.normal(0, 50, 20000).astype(np.int16)
    want, _ = ol.demod_batch("port", x)
    got = emu_lib.demod(x, [30000])
    check(want, got, "idle")
