"""Generator building blocks against the port (encode -> decode round trips).  CPU only."""
import ctypes

import numpy as np

import oracle_lib as ol
from m17_cxx_demod_b200 import gen


def p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def test_generator_crc_and_golay_agree_with_port():
    g, o = gen.lib(), ol.port()
    o.m17port_crc16.restype = ctypes.c_uint16; o.m17port_golay_encode24.restype = ctypes.c_uint32
    rng = np.random.default_rng(1)
    for n in (1, 9, 28, 30, 200):
        d = rng.integers(0, 256, n).astype(np.uint8)
        assert g.m17gen_crc16(p(d), ctypes.c_size_t(n)) == o.m17port_crc16(p(d), ctypes.c_size_t(n))
    for v in (0, 1, 0xD78, 0xFFF, 0x555):
        assert g.m17gen_golay24_encode(ctypes.c_uint16(v)) == o.m17port_golay_encode24(ctypes.c_uint16(v))
    cs = np.zeros(6, np.uint8)
    g.m17gen_encode_callsign(b"WX9O", p(cs))
    assert cs.tolist() == [0x00, 0x00, 0x00, 0x0f, 0x8a, 0xd7]      # tests/LinkSetupFrameTest.cpp


def decode_frames(frames, sync):
    fr = np.ascontiguousarray(frames, np.int8); sy = np.ascontiguousarray(sync, np.uint8)
    rec = np.zeros(64, ol.RECORD_DTYPE); n = len(fr)
    costs = np.zeros(n, np.int64); states = np.zeros(n, np.uint8); res = np.zeros(n, np.uint8)
    lib = ol.port(); lib.m17port_frame_decoder.restype = ctypes.c_size_t
    k = lib.m17port_frame_decoder(p(fr), p(sy), ctypes.c_size_t(n), p(rec), ctypes.c_size_t(64), p(costs), p(states), p(res))
    return rec[:k], costs, states


def soft(bits):
    return np.where(bits > 0, 7, -7).astype(np.int8)


def test_encode_decode_round_trip_all_frame_types():
    g = gen.lib()
    rng = np.random.default_rng(2)
    lsf = np.zeros(30, np.uint8); lsf[:6] = 0xFF
    g.m17gen_encode_callsign(b"AB1CDE", p(lsf[6:12])); lsf[13] = 5
    lsf[14:28] = rng.integers(0, 256, 14)
    crc = g.m17gen_crc16(p(lsf), ctypes.c_size_t(28)); lsf[28] = crc >> 8; lsf[29] = crc & 0xFF
    bits = np.zeros(368, np.uint8)
    frames, sync, payloads = [], [], []
    g.m17gen_lsf_frame_bits(p(lsf), p(bits)); frames.append(soft(bits)); sync.append(0)
    for f in range(7):
        pl = rng.integers(0, 256, 16).astype(np.uint8); payloads.append(pl)
        g.m17gen_stream_frame_bits(p(lsf), f % 6, f, p(pl), p(bits)); frames.append(soft(bits)); sync.append(1)
    rec, costs, states = decode_frames(frames, sync)
    assert rec[0]["type"] == 0 and rec[0]["lsf"].tolist() == lsf.tolist() and rec[0]["viterbi_cost"] == 0
    assert len(rec) == 8
    for f in range(7):
        r = rec[1 + f]
        assert r["type"] == 2 and r["viterbi_cost"] == 0
        assert r["payload"][:2].tolist() == [0, f] and r["payload"][2:18].tolist() == payloads[f].tolist()

    # LSF recovery from LICH: six stream frames with the decoder in LSF state
    frames, sync = [], []
    frames.append(soft(rng.integers(0, 2, 368).astype(np.uint8))); sync.append(0)       # garbage LSF -> CRC fail
    for f in range(6):
        pl = rng.integers(0, 256, 16).astype(np.uint8)
        g.m17gen_stream_frame_bits(p(lsf), f, f, p(pl), p(bits)); frames.append(soft(bits)); sync.append(1)
    rec, costs, states = decode_frames(frames, sync)
    assert [int(t) for t in rec["type"]] == [1] * 6 + [0]
    assert rec[-1]["lsf"].tolist() == lsf.tolist() and states[-1] == 1

    # packet mode (RAW -> BASIC_PACKET) and BERT
    lsf[13] = 2
    crc = g.m17gen_crc16(p(lsf), ctypes.c_size_t(28)); lsf[28] = crc >> 8; lsf[29] = crc & 0xFF
    frames, sync = [], []
    g.m17gen_lsf_frame_bits(p(lsf), p(bits)); frames.append(soft(bits)); sync.append(0)
    d26 = rng.integers(0, 256, 26).astype(np.uint8); d26[25] = 0x80 | (9 << 2)
    g.m17gen_packet_frame_bits(p(d26), p(bits)); frames.append(soft(bits)); sync.append(2)
    st = ctypes.c_uint16(1)
    g.m17gen_bert_frame_bits(ctypes.byref(st), p(bits)); frames.append(soft(bits)); sync.append(3)
    rec, costs, states = decode_frames(frames, sync)
    assert [int(t) for t in rec["type"]] == [0, 3, 5]
    assert rec[1]["payload"][:25].tolist() == d26[:25].tolist() and rec[1]["payload"][25] == d26[25] & 0xFC
    # PRBS9 from state 1 (tests/UtilTest.cpp:208-230): the first BERT bits follow the LFSR
    stt, want = 1, []
    for _ in range(197):
        r = ((stt >> 8) ^ (stt >> 4)) & 1; stt = ((stt << 1) | r) & 0x1FF; want.append(r)
    got = np.unpackbits(rec[2]["payload"][:25])[:197]
    assert got.tolist() == want


def test_generated_waveform_is_deterministic_and_bounded():
    ps = [gen.default_params(seed=5, snr_db=10.0, skew_ppm=250.0, dc_offset=100.0)]
    a = gen.generate(ps, 30000); b = gen.generate(ps, 30000)
    assert np.array_equal(a, b)
    assert a.dtype == np.int16 and np.abs(a.astype(int)).max() > 10000
    w = gen.workload_c2(n_streams=4, seconds=0.5)
    assert len(w.params) == 4 and w.n_samples == 24000
