"""The port against golden vectors produced by the UNMODIFIED reference headers
(tests/golden/make_golden.py -> ref_vectors.npz, ref_streams.npz).  CPU only; runs
anywhere because it reads only the committed fixtures."""
import ctypes
import os

import numpy as np
import pytest

import oracle_lib as ol

HERE = os.path.dirname(os.path.abspath(__file__))
V = np.load(os.path.join(HERE, "golden", "ref_vectors.npz"))
S = np.load(os.path.join(HERE, "golden", "ref_streams.npz"))


def p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def bits_equal(a, b):
    a = np.ascontiguousarray(a); b = np.ascontiguousarray(b)
    if a.dtype.kind == "f":
        return np.array_equal(a.view(np.uint32), b.view(np.uint32))
    return np.array_equal(a, b)


def test_scale_all_int16():
    lib = ol.port(); lib.m17port_scale.restype = ctypes.c_float
    got = np.array([lib.m17port_scale(ctypes.c_int16(int(s))) for s in range(-32768, 32768)], np.float32)
    assert bits_equal(got, V["scale_out"])


def test_fir_and_taps():
    lib = ol.port()
    taps = np.zeros(150, np.float32); lib.m17port_rrc_taps(p(taps))
    assert bits_equal(taps, V["taps"])
    x = V["fir_in"]; y = np.zeros_like(x)
    lib.m17port_fir(p(x), p(y), ctypes.c_size_t(len(x)))
    assert bits_equal(y, V["fir_out"])


def test_sdft_dcd():
    lib = ol.port()
    c = np.zeros(4, np.float32); lib.m17port_sdft_coeffs(p(c))
    assert bits_equal(c, V["sdft_coeffs"])
    x = V["sdft_in"]; o = np.zeros(4 * len(x), np.float32)
    lib.m17port_sdft(p(x), ctypes.c_size_t(len(x)), p(o))
    assert bits_equal(o, V["sdft_out"])
    lvl = np.zeros(64, np.float32); trg = np.zeros(64, np.uint8)
    lib.m17port_dcd.restype = ctypes.c_size_t
    k = lib.m17port_dcd(p(x), ctypes.c_size_t(len(x)), ctypes.c_size_t(384), p(lvl), p(trg))
    assert k == len(V["dcd_level"]) and bits_equal(lvl[:k], V["dcd_level"]) and bits_equal(trg[:k], V["dcd_trig"])
    assert V["dcd_trig"].any() and not V["dcd_trig"].all()


def test_limit_iir():
    x = V["iir_in"]; o = np.zeros_like(x)
    ol.port().m17port_limit_iir(p(x), p(o), ctypes.c_size_t(len(x)))
    assert bits_equal(o, V["iir_out"])


def test_correlator():
    cin, cout = V["corr_in"], V["corr_out"]
    i = j = 0
    lib = ol.port()
    while i < len(cin):
        n, si = int(cin[i]), int(cin[i + 1])
        x = np.ascontiguousarray(cin[i + 2:i + 2 + n])
        corr = np.zeros(4, np.float32); mm = np.zeros(2, np.float32); lim = np.zeros(1, np.float32); idx = np.zeros(1, np.uint8)
        lib.m17port_correlator(p(x), ctypes.c_size_t(n), ctypes.c_uint8(si), p(corr), p(mm), p(lim), p(idx))
        got = np.concatenate([corr, mm, lim, idx.astype(np.float32)])
        assert bits_equal(got, cout[j:j + 8])
        i += 2 + n; j += 8
    assert j == len(cout)


def test_clock_recovery_kalman():
    n = len(V["clk_ops"])
    si = np.zeros(n, np.uint8); ce = np.zeros(n, np.float32); se = np.zeros(n, np.float32)
    ol.port().m17port_clock(p(V["clk_ops"]), p(V["clk_args"]), p(V["clk_dt"]), ctypes.c_size_t(n), p(si), p(ce), p(se))
    assert bits_equal(si, V["clk_si"]) and bits_equal(ce, V["clk_ce"]) and bits_equal(se, V["clk_se"])


def test_freqdev_kalman_incl_nan_reset():
    n = len(V["dev_mn"])
    off = np.zeros(n, np.float32); idev = np.zeros(n, np.float32)
    ol.port().m17port_freqdev(p(V["dev_mn"]), p(V["dev_mx"]), p(V["dev_rs"]), ctypes.c_size_t(n), p(off), p(idev))
    assert bits_equal(off, V["dev_off"]) and bits_equal(idev, V["dev_idev"])


def test_evm():
    x = V["evm_in"]; o = np.zeros_like(x)
    ol.port().m17port_evm(p(x), ctypes.c_size_t(len(x)), p(o))
    assert bits_equal(o, V["evm_out"])


def test_llr():
    lib = ol.port()
    x = V["llr_in"]; o = np.zeros((len(x), 2), np.int8)
    for i, s in enumerate(x):
        lib.m17port_llr(ctypes.c_float(float(s)), p(o[i]))
    assert np.array_equal(o, V["llr_out"])
    thr = np.zeros(43, np.float32); pairs = np.zeros(86, np.int8)
    lib.m17port_llr_table(p(thr), p(pairs))
    assert bits_equal(thr, V["llr_thr"]) and np.array_equal(pairs, V["llr_pairs"])


def test_crc_golay():
    lib = ol.port(); lib.m17port_crc16.restype = ctypes.c_uint16; lib.m17port_golay_encode24.restype = ctypes.c_uint32
    data = V["crc_in"]
    got = np.array([lib.m17port_crc16(p(np.ascontiguousarray(data[30 * i:30 * i + 30])), ctypes.c_size_t(30)) for i in range(64)], np.uint16)
    assert np.array_equal(got, V["crc_out"])
    for w, want, ok in zip(V["golay_in"], V["golay_out"], V["golay_ok"]):
        o = ctypes.c_uint32()
        r = lib.m17port_golay_decode(ctypes.c_uint32(int(w)), ctypes.byref(o))
        assert bool(r) == bool(ok) and o.value == int(want)
    enc = np.array([lib.m17port_golay_encode24(ctypes.c_uint16(i)) for i in range(4096)], np.uint32)
    assert np.array_equal(enc, V["golay_enc"])
    assert 0 < V["golay_ok"].sum() < len(V["golay_ok"])


@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_viterbi_random_soft_input(kind):
    lib = ol.port(); lib.m17port_viterbi.restype = ctypes.c_size_t
    ins, outs, costs = V[f"vit{kind}_in"], V[f"vit{kind}_out"], V[f"vit{kind}_cost"]
    for x, want, c in zip(ins, outs, costs):
        o = np.zeros(len(want), np.uint8)
        got = lib.m17port_viterbi(kind, p(np.ascontiguousarray(x)), p(o))
        assert got == c and np.array_equal(o, want)


def test_tables():
    tabs = [np.zeros(368, np.uint16), np.zeros(368, np.int8), np.zeros(61, np.int8), np.zeros(12, np.int8), np.zeros(8, np.int8)]
    ol.port().m17port_tables(*[p(t) for t in tabs])
    for t, k in zip(tabs, ("tab_il", "tab_dc", "tab_p1", "tab_p2", "tab_p3")):
        assert np.array_equal(t, V[k])


def test_frame_decoder_state_machine():
    fr, sy = V["fd_frames"], V["fd_sync"]
    rec = np.zeros(128, ol.RECORD_DTYPE); n = len(fr)
    costs = np.zeros(n, np.int64); states = np.zeros(n, np.uint8); res = np.zeros(n, np.uint8)
    lib = ol.port(); lib.m17port_frame_decoder.restype = ctypes.c_size_t
    k = lib.m17port_frame_decoder(p(fr), p(sy), ctypes.c_size_t(n), p(rec), ctypes.c_size_t(128), p(costs), p(states), p(res))
    want = V["fd_rec"].view(ol.RECORD_DTYPE)
    assert k == len(want)
    assert ol.records_equal(rec[:k], want), ol.describe_mismatch(rec[:k], want)
    assert np.array_equal(costs, V["fd_costs"]) and np.array_equal(states, V["fd_states"]) and np.array_equal(res, V["fd_results"])
    # the corpus exercises every frame type
    assert set(want["type"].tolist()) == {0, 1, 2, 3, 5} or set(want["type"].tolist()) >= {0, 1, 2, 5}


def test_whole_demodulator_streams():
    x = S["x"]
    step = int(S["trace_step"][0])
    rec, tr = ol.demod_batch("port", x, trace=True)
    types = set()
    for s in range(x.shape[0]):
        want = S[f"rec{s}"].view(ol.RECORD_DTYPE)
        assert ol.records_equal(rec[s], want), f"stream {s}: " + ol.describe_mismatch(rec[s], want)
        wt = S[f"trace{s}"].view(ol.TRACE_DTYPE)
        assert np.array_equal(np.ascontiguousarray(tr[s][::step]).view(np.uint8), wt.view(np.uint8)), f"stream {s}: internal state trace differs"
        types |= set(want["type"].tolist())
    assert types == {0, 1, 2, 3, 4, 5}        # LSF, LICH, STREAM, BASIC_PACKET, FULL_PACKET, BERT all covered


def test_cli_quirks_invert_and_eof_repeat():
    # apps/m17-demod.cpp:484-490: -i and the last sample processed twice
    rec, _ = ol.demod_batch("port", S["x"][:1], flags=1 | 2)
    want = S["rec_invert_eof"].view(ol.RECORD_DTYPE)
    assert ol.records_equal(rec[0], want)
