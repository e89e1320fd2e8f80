"""Generate the committed golden fixtures.  Run ONCE in the build container, where
/root/reference and oracle/_ref exist:   python tests/golden/make_golden.py

  ref_kat.json      vectors copied out of the reference's own unit tests
                    (tests/ViterbiTest.cpp:173-284, tests/ConvolutionTest.cpp:35-65)
  ref_vectors.npz   inputs + outputs of individual reference blocks, produced by the
                    UNMODIFIED reference headers through oracle/_ref/libm17ref.so
  ref_streams.npz   whole-demodulator cases: int16 baseband, the reference's frame
                    records and a decimated internal-state trace

Nothing at test time reads /root/reference; the tests only read these files.
"""
import ctypes
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as ol  # noqa: E402
from m17_cxx_demod_b200 import gen  # noqa: E402

REF = "/root/reference"


def p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def parse_array(src, name, occurrence=0):
    m = [x for x in re.finditer(r"std::array<[^>]+>\s+" + name + r"\s*=\s*\{([^}]*)\}", src)]
    return [int(v) for v in m[occurrence].group(1).replace("\n", " ").split(",") if v.strip()]


def kat():
    vt = open(os.path.join(REF, "tests", "ViterbiTest.cpp")).read()
    ct = open(os.path.join(REF, "tests", "ConvolutionTest.cpp")).read()
    out = {
        "source": "mobilinkd/m17-cxx-demod tests/ViterbiTest.cpp, tests/ConvolutionTest.cpp",
        "conv_input8": parse_array(ct, "input"),
        "conv_expected24": parse_array(ct, "expected"),
        "lsf_expected240": parse_array(vt[vt.index("decode_ber_lsf"):], "expected"),
        "lsf_encoded488": parse_array(vt[vt.index("decode_ber_lsf"):], "encoded"),
    }
    assert len(out["lsf_expected240"]) == 240 and len(out["lsf_encoded488"]) == 488
    json.dump(out, open(os.path.join(HERE, "ref_kat.json"), "w"))


def vectors():
    ref = ol.ref()
    rng = np.random.default_rng(20261104)
    v = {}
    # scaling: all 65536 int16 values
    ref.m17ref_scale.restype = ctypes.c_float
    allv = np.arange(-32768, 32768, dtype=np.int32)
    v["scale_out"] = np.array([ref.m17ref_scale(ctypes.c_int16(int(s))) for s in allv], dtype=np.float32)
    # FIR
    x = (rng.integers(-30000, 30000, 3000).astype(np.float64) / 41067.0).astype(np.float32)
    y = np.zeros_like(x)
    ref.m17ref_fir(p(x), p(y), ctypes.c_size_t(len(x)))
    v["fir_in"], v["fir_out"] = x, y
    taps = np.zeros(150, np.float32); ref.m17ref_rrc_taps(p(taps)); v["taps"] = taps
    # SDFT + DCD
    x = (rng.normal(0, 0.2, 6000) + np.sin(np.arange(6000) * 2 * np.pi * 2400 / 48000) * (np.arange(6000) > 2500)).astype(np.float32)
    o = np.zeros(4 * len(x), np.float32); ref.m17ref_sdft(p(x), ctypes.c_size_t(len(x)), p(o))
    v["sdft_in"], v["sdft_out"] = x, o
    lvl = np.zeros(64, np.float32); trg = np.zeros(64, np.uint8)
    ref.m17ref_dcd.restype = ctypes.c_size_t
    k = ref.m17ref_dcd(p(x), ctypes.c_size_t(len(x)), ctypes.c_size_t(384), p(lvl), p(trg))
    v["dcd_level"], v["dcd_trig"] = lvl[:k], trg[:k]
    c = np.zeros(4, np.float32); ref.m17ref_sdft_coeffs(p(c)); v["sdft_coeffs"] = c
    # limit IIR
    x = rng.normal(0, 1.5, 4000).astype(np.float32); o = np.zeros_like(x)
    ref.m17ref_limit_iir(p(x), p(o), ctypes.c_size_t(len(x)))
    v["iir_in"], v["iir_out"] = x, o
    # correlator: several lengths / sample indices
    cin, cout = [], []
    for n, si in ((80, 0), (95, 3), (237, 7), (1000, 9), (83, 5)):
        x = rng.normal(0, 2.0, n).astype(np.float32)
        corr = np.zeros(4, np.float32); mm = np.zeros(2, np.float32); lim = np.zeros(1, np.float32); idx = np.zeros(1, np.uint8)
        ref.m17ref_correlator(p(x), ctypes.c_size_t(n), ctypes.c_uint8(si), p(corr), p(mm), p(lim), p(idx))
        cin.append(np.concatenate([[n, si], x]).astype(np.float32))
        cout.append(np.concatenate([corr, mm, lim, idx.astype(np.float32)]))
    v["corr_in"] = np.concatenate(cin); v["corr_out"] = np.concatenate(cout)
    # clock recovery / Kalman
    n = 600
    ops = rng.choice([1, 1, 1, 2, 2, 0], n).astype(np.uint8); ops[0] = 0
    args = rng.integers(0, 10, n).astype(np.uint8)
    dt = np.where(ops == 2, rng.integers(1, 12, n), rng.integers(1900, 1941, n)).astype(np.uint32)
    si = np.zeros(n, np.uint8); ce = np.zeros(n, np.float32); se = np.zeros(n, np.float32)
    ref.m17ref_clock(p(ops), p(args), p(dt), ctypes.c_size_t(n), p(si), p(ce), p(se))
    v["clk_ops"], v["clk_args"], v["clk_dt"], v["clk_si"], v["clk_ce"], v["clk_se"] = ops, args, dt, si, ce, se
    # frequency/deviation estimator (incl. a NaN-producing degenerate pair and resets)
    n = 500
    mn = (-3 + rng.normal(0, 0.3, n)).astype(np.float32) / np.float32(6); mx = (3 + rng.normal(0, 0.3, n)).astype(np.float32) / np.float32(6)
    mn[100] = mx[100] = 0.25; mn[101] = np.float32("nan")
    rs = (rng.random(n) < 0.03).astype(np.uint8)
    off = np.zeros(n, np.float32); idev = np.zeros(n, np.float32)
    ref.m17ref_freqdev(p(mn), p(mx), p(rs), ctypes.c_size_t(n), p(off), p(idev))
    v["dev_mn"], v["dev_mx"], v["dev_rs"], v["dev_off"], v["dev_idev"] = mn, mx, rs, off, idev
    # EVM
    x = rng.normal(0, 2, 1000).astype(np.float32); o = np.zeros_like(x)
    ref.m17ref_evm(p(x), ctypes.c_size_t(len(x)), p(o)); v["evm_in"], v["evm_out"] = x, o
    # LLR
    x = np.concatenate([np.linspace(-4, 4, 4001), rng.normal(0, 2, 2000)]).astype(np.float32)
    o = np.zeros((len(x), 2), np.int8)
    for i, s in enumerate(x):
        ref.m17ref_llr(ctypes.c_float(float(s)), p(o[i]))
    v["llr_in"], v["llr_out"] = x, o
    thr = np.zeros(43, np.float32); pairs = np.zeros(86, np.int8); ref.m17ref_llr_table(p(thr), p(pairs))
    v["llr_thr"], v["llr_pairs"] = thr, pairs
    # CRC16
    data = rng.integers(0, 256, 64 * 30).astype(np.uint8); ref.m17ref_crc16.restype = ctypes.c_uint16
    v["crc_in"] = data
    v["crc_out"] = np.array([ref.m17ref_crc16(p(data[30 * i:30 * i + 30]), ctypes.c_size_t(30)) for i in range(64)], np.uint16)
    # Golay: every weight <= 4 style corruption of random codewords + random words
    words = rng.integers(0, 1 << 24, 3000).astype(np.uint32)
    outw = np.zeros_like(words); ok = np.zeros(len(words), np.uint8)
    for i, w in enumerate(words):
        o32 = ctypes.c_uint32(); ok[i] = ref.m17ref_golay_decode(ctypes.c_uint32(int(w)), ctypes.byref(o32)); outw[i] = o32.value
    v["golay_in"], v["golay_out"], v["golay_ok"] = words, outw, ok
    ref.m17ref_golay_encode24.restype = ctypes.c_uint32
    v["golay_enc"] = np.array([ref.m17ref_golay_encode24(ctypes.c_uint16(i)) for i in range(4096)], np.uint32)
    # Viterbi on random soft input for each frame kind (0 = erasure allowed)
    ref.m17ref_viterbi.restype = ctypes.c_size_t
    for kind, (n_in, n_out) in enumerate(((488, 240), (296, 144), (420, 206), (402, 197))):
        ins, outs, costs = [], [], []
        for _ in range(12):
            x = rng.integers(-7, 8, n_in).astype(np.int8)
            o = np.zeros(n_out, np.uint8)
            costs.append(ref.m17ref_viterbi(kind, p(x), p(o))); ins.append(x); outs.append(o)
        v[f"vit{kind}_in"], v[f"vit{kind}_out"], v[f"vit{kind}_cost"] = np.array(ins), np.array(outs), np.array(costs, np.int64)
    # tables
    il = np.zeros(368, np.uint16); dc = np.zeros(368, np.int8); p1 = np.zeros(61, np.int8); p2 = np.zeros(12, np.int8); p3 = np.zeros(8, np.int8)
    ref.m17ref_tables(p(il), p(dc), p(p1), p(p2), p(p3))
    v["tab_il"], v["tab_dc"], v["tab_p1"], v["tab_p2"], v["tab_p3"] = il, dc, p1, p2, p3
    # frame decoder driven directly: valid frames from the generator (hard +-7) mixed with noise frames
    g = gen.lib()
    lsf = np.zeros(30, np.uint8); lsf[:6] = 0xFF
    g.m17gen_encode_callsign(b"N0CALL", p(lsf[6:12])); lsf[13] = 5
    crc = g.m17gen_crc16(p(lsf), ctypes.c_size_t(28)); lsf[28] = crc >> 8; lsf[29] = crc & 0xFF
    frames, sync = [], []

    def soft(bits, flips=0):
        s = (1 - 2 * bits.astype(np.int8)) * 0 + np.where(bits > 0, 7, -7).astype(np.int8)
        idx = rng.choice(368, flips, replace=False)
        s[idx] = -s[idx]
        s[rng.choice(368, 20, replace=False)] //= 3
        return s.astype(np.int8)

    bits = np.zeros(368, np.uint8)
    g.m17gen_lsf_frame_bits(p(lsf), p(bits)); frames.append(soft(bits, 3)); sync.append(0)
    for f in range(9):
        payload = rng.integers(0, 256, 16).astype(np.uint8)
        g.m17gen_stream_frame_bits(p(lsf), f % 6, f, p(payload), p(bits)); frames.append(soft(bits, f)); sync.append(1)
    # packet LSF then packet frames
    lsf2 = lsf.copy(); lsf2[13] = 4
    crc = g.m17gen_crc16(p(lsf2), ctypes.c_size_t(28)); lsf2[28] = crc >> 8; lsf2[29] = crc & 0xFF
    g.m17gen_lsf_frame_bits(p(lsf2), p(bits)); frames.append(soft(bits, 2)); sync.append(0)
    for f in range(3):
        d26 = rng.integers(0, 256, 26).astype(np.uint8); d26[25] = (f << 2) if f < 2 else (0x80 | (17 << 2))
        g.m17gen_packet_frame_bits(p(d26), p(bits)); frames.append(soft(bits, 4)); sync.append(2)
    st = ctypes.c_uint16(1)
    for f in range(3):
        g.m17gen_bert_frame_bits(ctypes.byref(st), p(bits)); frames.append(soft(bits, 5)); sync.append(3)
    # LICH recovery: stream frames while the decoder is in LSF state, then noise, wrong-state sync types
    frames.append(rng.integers(-7, 8, 368).astype(np.int8)); sync.append(0)      # bad LSF -> CRC fail
    for f in range(8):
        payload = rng.integers(0, 256, 16).astype(np.uint8)
        g.m17gen_stream_frame_bits(p(lsf), f % 6, 100 + f, p(payload), p(bits)); frames.append(soft(bits, 2)); sync.append(1)
    frames.append(rng.integers(-7, 8, 368).astype(np.int8)); sync.append(2)
    frames.append(rng.integers(-7, 8, 368).astype(np.int8)); sync.append(1)
    frames.append(rng.integers(-7, 8, 368).astype(np.int8)); sync.append(3)
    frames.append(rng.integers(-7, 8, 368).astype(np.int8)); sync.append(1)
    fr = np.array(frames, np.int8); sy = np.array(sync, np.uint8)
    rec = np.zeros(128, ol.RECORD_DTYPE); costs = np.zeros(len(fr), np.int64); states = np.zeros(len(fr), np.uint8); res = np.zeros(len(fr), np.uint8)
    ref.m17ref_frame_decoder.restype = ctypes.c_size_t
    k = ref.m17ref_frame_decoder(p(fr), p(sy), ctypes.c_size_t(len(fr)), p(rec), ctypes.c_size_t(128), p(costs), p(states), p(res))
    v["fd_frames"], v["fd_sync"], v["fd_rec"], v["fd_costs"], v["fd_states"], v["fd_results"] = fr, sy, rec[:k].view(np.uint8), costs, states, res
    np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **v)


def streams():
    cases = [
        dict(seed=51, mode=gen.STREAM, snr_db=3.0, skew_ppm=-100.0, dc_offset=100.0, lead_in=4000, n_frames=20),
        dict(seed=52, mode=gen.PACKET, snr_db=3.0, skew_ppm=0.0, dc_offset=200.0, lead_in=4000, n_frames=20),
        dict(seed=51, mode=gen.PACKET, snr_db=3.0, skew_ppm=-100.0, dc_offset=100.0, lead_in=4000, n_frames=20, packet_encapsulated=1),
        dict(seed=52, mode=gen.BERT, snr_db=10.0, skew_ppm=0.0, dc_offset=200.0, lead_in=1000, n_frames=20),
        dict(seed=51, mode=gen.STREAM, snr_db=10.0, skew_ppm=-100.0, dc_offset=100.0, lead_in=0, n_frames=20),
    ]
    n = 72000
    x = gen.generate([gen.default_params(**c) for c in cases], n)
    rec, tr = ol.demod_batch("ref", x, trace=True)
    out = {"x": x, "trace_step": np.array([37])}
    for s in range(len(cases)):
        out[f"rec{s}"] = rec[s].view(np.uint8)
        out[f"trace{s}"] = tr[s][::37].copy().view(np.uint8)
    # CLI quirk case: EOF repeat + invert
    rec_i, _ = ol.demod_batch("ref", x[:1], flags=1 | 2)
    out["rec_invert_eof"] = rec_i[0].view(np.uint8)
    np.savez_compressed(os.path.join(HERE, "ref_streams.npz"), **out)
    print("records per case:", [len(r) for r in rec], len(rec_i[0]))


if __name__ == "__main__":
    assert ol.have_ref(), "build oracle/_ref first (make -C oracle ref)"
    kat()
    vectors()
    streams()
    for f in ("ref_kat.json", "ref_vectors.npz", "ref_streams.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
