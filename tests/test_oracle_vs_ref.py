"""Live cross-check: oracle/m17_port.c against oracle/_ref (the unmodified reference
headers) on freshly generated streams, record for record and state for state.  Needs
oracle/_ref/libm17ref.so, which exists wherever build() ran with /root/reference present
(and travels to the GPU box); skipped otherwise."""
import numpy as np
import pytest

import oracle_lib as ol
from m17_cxx_demod_b200 import gen

pytestmark = pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (no /root/reference here)")


def random_params(rng, mode, n):
    return [gen.default_params(seed=int(rng.integers(1, 1 << 40)), mode=mode, snr_db=float(rng.uniform(-2, 30)),
                               skew_ppm=float(rng.uniform(-500, 500)), dc_offset=float(rng.uniform(-400, 400)), repeat=1,
                               gap=int(rng.integers(2000, 9000)), lead_in=int(rng.integers(0, 9600)),
                               n_frames=int(rng.integers(0, 30)), packet_encapsulated=int(rng.integers(0, 2)))
            for _ in range(n)]


@pytest.mark.parametrize("mode", [gen.STREAM, gen.PACKET, gen.BERT, gen.MIXED])
def test_port_equals_reference(mode):
    rng = np.random.default_rng(100 + mode)
    x = gen.generate(random_params(rng, mode, 12), 100000)
    ref_rec, ref_tr = ol.demod_batch("ref", x, trace=True)
    port_rec, port_tr = ol.demod_batch("port", x, trace=True)
    n_rec = 0
    for s in range(x.shape[0]):
        assert ol.records_equal(ref_rec[s], port_rec[s]), f"stream {s}: " + ol.describe_mismatch(ref_rec[s], port_rec[s])
        assert np.array_equal(ref_tr[s].view(np.uint8), port_tr[s].view(np.uint8)), f"stream {s}: state trace differs"
        n_rec += len(ref_rec[s])
    assert n_rec > 50


def test_port_equals_reference_invert_and_eof():
    rng = np.random.default_rng(7)
    ps = random_params(rng, gen.STREAM, 4)
    for q in ps:
        q.invert = 1
    x = gen.generate(ps, 80000)
    a, _ = ol.demod_batch("ref", x, flags=3)
    b, _ = ol.demod_batch("port", x, flags=3)
    for s in range(4):
        assert ol.records_equal(a[s], b[s])


def test_all_zero_input_never_decodes():
    # DataCarrierDetect.h:65: 0/0 = NaN poisons the DCD level; nothing is ever decoded
    x = np.zeros((1, 20000), np.int16)
    a, ta = ol.demod_batch("ref", x, trace=True)
    b, tb = ol.demod_batch("port", x, trace=True)
    assert len(a[0]) == 0 and len(b[0]) == 0
    assert np.isnan(ta[0]["dcd_level"][-1]) and np.isnan(tb[0]["dcd_level"][-1])


def test_reference_cli_binary_matches_library(tmp_path):
    import subprocess
    x = gen.generate([gen.default_params(seed=47, snr_db=3.0, lead_in=4000, n_frames=20, skew_ppm=-100.0, dc_offset=100.0)], 72000)
    f = tmp_path / "x.s16"
    x[0].tofile(f)
    out = subprocess.run([ol.ref_binary()], stdin=open(f, "rb"), capture_output=True, check=True).stdout.decode().splitlines()
    rec, _ = ol.demod_batch("ref", x, flags=2)
    assert len(out) == len(rec[0]) > 10
    for line, r in zip(out, rec[0]):
        assert f"cost={int(r['viterbi_cost'])} " in line and ol.TYPE_NAMES[int(r["type"])] in line


def test_correlate_trace_port_equals_reference():
    """Correlator.h:43-64 per sample (limit_ and the four sync-word correlations): the restatement against the
    reference's own Correlator class -- the oracle of the GPU correlator tap (test_gpu_parity.py)."""
    import ctypes
    rng = np.random.default_rng(5)
    x = rng.normal(0, 1.5, 4000).astype(np.float32)
    a = np.zeros((4000, 5), np.float32); b = np.zeros((4000, 5), np.float32)
    ol.ref().m17ref_correlate_trace(x.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(4000), a.ctypes.data_as(ctypes.c_void_p))
    ol.port().m17port_correlate_trace(x.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(4000), b.ctypes.data_as(ctypes.c_void_p))
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
