"""m17b_filter_noalloc (no counterpart in the reference): the matched
filter through the C ABI without any allocation -- its scratch is a static __device__ array -- against the existing
entry point m17b_debug_filter, bit for bit (the exact filter mode is bit-identical by construction, the tolerance all
filter tests here use), on two calls in a row that share the scratch.  Collected last on purpose (pytest -x)."""
import ctypes

import numpy as np
import pytest

import m17_cxx_demod_b200 as m
from m17_cxx_demod_b200 import gen

pytestmark = pytest.mark.gpu

OK, EINVAL, ETOOBIG = 0, -1, -7
MAX_STREAMS, MAX_SAMPLES = 128, 4000


def _noalloc(lib, x_dev, n_streams, n, stride, out_dev, flags=0, stream=None):
    return lib.m17b_filter_noalloc(0, ctypes.c_void_p(x_dev.data_ptr()), n_streams, n, stride, ctypes.c_void_p(out_dev.data_ptr()),
                                   flags, ctypes.c_void_p(stream))


def test_filter_noalloc_matches_debug_filter_on_two_calls_in_a_row():
    import torch
    lib = m.load_library()
    rng = np.random.default_rng(123)
    ps = [gen.default_params(seed=9000 + k, mode=k % 4, snr_db=5.0 + k % 20, lead_in=100 * (k % 7), n_frames=2, repeat=1) for k in range(MAX_STREAMS)]
    # first call: the largest size the entry point takes (row stride wider than the rows); second call: a smaller,
    # ragged one with other data, which must not see anything the first left in the shared scratch
    x1 = np.ascontiguousarray(gen.generate(ps, 12000)[:, 2000:2000 + MAX_SAMPLES])      # preamble / LSF of a transmission
    x2 = (rng.integers(-32768, 32768, (37, 1003))).astype(np.int16)
    cases = [(x1, MAX_SAMPLES + 24), (x2, 1003)]
    got = []
    # the new entry point runs before any context exists in this process: it sets itself up (tables, scratch address)
    for x, stride in cases:
        ns, n = x.shape
        xd = torch.zeros((ns, stride), dtype=torch.int16, device="cuda")
        xd[:, :n] = torch.from_numpy(x).cuda()
        yd = torch.full((ns, n), float("nan"), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        assert _noalloc(lib, xd, ns, n, stride, yd) == OK
        torch.cuda.synchronize()
        got.append(yd.cpu().numpy())
    for (x, _), y in zip(cases, got):
        with m.Demodulator(x.shape[0], x.shape[1]) as d:
            want, _ = d.debug_filter(x)
        assert np.array_equal(y.view(np.uint32), want.view(np.uint32)), f"{int((y.view(np.uint32) != want.view(np.uint32)).sum())} outputs differ"
    assert np.abs(got[0]).max() > 0.5

    # FMA mode and the inverted input go through the same flags as the existing entry point's contexts
    ns, n = x2.shape
    xd = torch.from_numpy(x2).cuda()
    yd = torch.empty((ns, n), dtype=torch.float32, device="cuda")
    assert _noalloc(lib, xd, ns, n, n, yd, flags=m.FIR_FMA | m.INVERT) == OK
    torch.cuda.synchronize()
    with m.Demodulator(ns, n, flags=m.FIR_FMA | m.INVERT) as d:
        want, _ = d.debug_filter(x2)
    assert np.array_equal(yd.cpu().numpy().view(np.uint32), want.view(np.uint32))

    # sizes beyond the static scratch are refused with their own error, nothing is launched
    big = torch.zeros((MAX_STREAMS + 1, 64), dtype=torch.int16, device="cuda")
    out = torch.zeros((MAX_STREAMS + 1, MAX_SAMPLES + 1), dtype=torch.float32, device="cuda")
    assert _noalloc(lib, big, MAX_STREAMS + 1, 64, 64, out) == ETOOBIG
    assert _noalloc(lib, big, 1, MAX_SAMPLES + 1, MAX_SAMPLES + 1, out) == ETOOBIG
    assert b"static scratch" in lib.m17b_strerror(ETOOBIG)
    assert _noalloc(lib, big, 4, 64, 32, out) == EINVAL               # stride shorter than a row
