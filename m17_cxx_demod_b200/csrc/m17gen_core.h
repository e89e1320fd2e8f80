// m17gen_core.h -- the synthetic M17 baseband generator's logic, shared by its two builds:
//   csrc/m17gen.cpp     the CPU generator (libm17gen.so, one host thread per stream), and
//   csrc/m17gen_dev.cu  the device-side modulator (libm17gen_cuda.so): frame construction, RRC pulse shaping,
//                       resampling and channel impairments on the GPU, so that every channel of a large run can
//                       be its own waveform without generating and uploading it from the host.
// Everything here is plain integer and IEEE double arithmetic (no libm calls whose results could differ between
// glibc and CUDA: the one logarithm the Gaussian noise needs is computed by log_pos() below on both sides; the
// host build uses -ffp-contract=off, the device build -fmad=false), so the two builds give identical int16
// samples for identical parameters -- tests/test_gpu_parity.py::test_device_modulator_* checks exactly that.
//
// Restates, with random payloads, what the reference's transmit CLI does:
//   symbols_to_baseband   apps/m17-mod.cpp:204-224  (150-tap RRC, x7168, 10 sps)
//   send_preamble         apps/m17-mod.cpp:264-280  (48 bytes of 0x77)
//   sync words            apps/m17-mod.cpp:282-287  (+ packet 0x75FF from M17Demodulator.h:156)
//   output_eot            apps/m17-mod.cpp:289-308
//   send_lsf              apps/m17-mod.cpp:310-386
//   make_data_frame       apps/m17-mod.cpp:407-440
//   make_bert_frame       apps/m17-mod.cpp:442-504
//   make_lich_segment     apps/m17-mod.cpp:509-548
//   send_audio_frame      apps/m17-mod.cpp:550-564
// Channel impairments (AWGN, clock skew, DC offset, gain) are this project's own.
#pragma once

#include "m17gen.h"

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define M17G_FN __host__ __device__ inline
#else
#define M17G_FN inline
#endif

namespace m17gen {

// Root-raised-cosine taps (alpha 0.5, 10 samples/symbol, 150 taps incl. the trailing zero): the same design the
// reference ships (apps/m17-mod.cpp:36-38); only the first 75 are stored, the filter is symmetric about tap 74.
#define M17G_RRC_HALF 0.0029364388513841593, 0.0031468394550958484, 0.002699564567597445, 0.001661182944400927, \
    0.00023319405581230247, -0.0012851320781224025, -0.0025577136087664687, -0.0032843366522956313, \
    -0.0032697038088887226, -0.0024733964729590865, -0.0010285696910973807, 0.0007766690889758685, \
    0.002553421969211845, 0.0038920145144327816, 0.004451886520053017, 0.00404219185231544, \
    0.002674727068399207, 0.0005756567993179152, -0.0018493784971116507, -0.004092346891623224, \
    -0.005648131453822014, -0.006126925416243605, -0.005349511529163396, -0.003403189203405097, \
    -0.0006430502751187517, 0.002365929161655135, 0.004957956568090113, 0.006506845894531803, \
    0.006569574194782443, 0.0050017573119839134, 0.002017321931508163, -0.0018256054303579805, \
    -0.00571615173291049, -0.008746639552588416, -0.010105075751866371, -0.009265784007800534, \
    -0.006136551625729697, -0.001125978562075172, 0.004891777252042491, 0.01071805138282269, \
    0.01505751553351295, 0.01679337935001369, 0.015256245142156299, 0.01042830577908502, \
    0.003031522725559901, -0.0055333532968188165, -0.013403099825723372, -0.018598682349642525, \
    -0.01944761739590459, -0.015005271935951746, -0.0053887880354343935, 0.008056525910253532, \
    0.022816244158307273, 0.035513467692208076, 0.04244131815783876, 0.04025481153629372, \
    0.02671818654865632, 0.0013810216516704976, -0.03394615682795165, -0.07502635967975885, \
    -0.11540977897637611, -0.14703962203941534, -0.16119995609538576, -0.14969512896336504, \
    -0.10610329539459686, -0.026921412469634916, 0.08757875030779196, 0.23293327870303457, \
    0.4006012210123992, 0.5786324696325503, 0.7528286479934068, 0.908262741447522, \
    1.0309661131633199, 1.1095611856548013, 1.1366197723675815
// M17 decorrelation sequence (M17Randomizer.h:16-22; M17 spec table).
#define M17G_DC \
    0xd6, 0xb5, 0xe2, 0x30, 0x82, 0xFF, 0x84, 0x62, 0xba, 0x4e, 0x96, 0x90, 0xd8, 0x98, 0xdd, 0x5d, \
    0x0c, 0xc8, 0x52, 0x43, 0x91, 0x1d, 0xf8, 0x6e, 0x68, 0x2F, 0x35, 0xda, 0x14, 0xea, 0xcd, 0x76, \
    0x19, 0x8d, 0xd5, 0x80, 0xd1, 0x33, 0x87, 0x13, 0x57, 0x18, 0x2d, 0x29, 0x78, 0xc3

static const double h_rrc_half[75] = {M17G_RRC_HALF};
static const uint8_t h_dc[46] = {M17G_DC};
#ifdef __CUDACC__
static __device__ const double d_rrc_half[75] = {M17G_RRC_HALF};
static __device__ const uint8_t d_dc[46] = {M17G_DC};
#endif

M17G_FN double rrc_half(int i)
{
#ifdef __CUDA_ARCH__
    return d_rrc_half[i];
#else
    return h_rrc_half[i];
#endif
}

M17G_FN uint8_t dc_byte(int i)
{
#ifdef __CUDA_ARCH__
    return d_dc[i];
#else
    return h_dc[i];
#endif
}

M17G_FN double rrc_tap(int i)   // i in [0,150)
{
    if (i < 0 || i >= 149) return 0.0;
    return i <= 74 ? rrc_half(i) : rrc_half(148 - i);
}

M17G_FN int popcount32(uint32_t v)
{
    v = v - ((v >> 1) & 0x55555555u);
    v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
    return (int) ((((v + (v >> 4)) & 0x0F0F0F0Fu) * 0x01010101u) >> 24);
}

// splitmix64 / xoshiro256**: payloads, call signs, traffic mix, start phase
M17G_FN uint64_t splitmix64(uint64_t& z)
{
    z += 0x9E3779B97F4A7C15ull;
    uint64_t y = z;
    y = (y ^ (y >> 30)) * 0xBF58476D1CE4E5B9ull;
    y = (y ^ (y >> 27)) * 0x94D049BB133111EBull;
    return y ^ (y >> 31);
}

struct Rng {
    uint64_t s[4];
    M17G_FN explicit Rng(uint64_t seed)
    {
        uint64_t z = seed;
        for (int i = 0; i != 4; ++i) s[i] = splitmix64(z);
    }
    M17G_FN static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    M17G_FN uint64_t next()
    {
        uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    M17G_FN double uniform() { return (double) (next() >> 11) * (1.0 / 9007199254740992.0); }
    M17G_FN uint32_t below(uint32_t n) { return uint32_t((next() >> 32) * uint64_t(n) >> 32); }
};

// ln(x) for 0 < x < 1 from +, -, *, / only (the same sequence of IEEE operations on the host and on the device):
// x = m 2^e with m in [sqrt(1/2), sqrt(2)), ln m = 2 atanh(s), s = (m - 1) / (m + 1), |s| < 0.1716, odd series
// up to s^21 (truncation below 3e-18).  Accurate to a few ulp, which is all a noise generator needs.
M17G_FN double log_pos(double x)
{
    int e = 0;
    while (x < 0.70710678118654752) { x *= 2.0; --e; }
    while (x >= 1.41421356237309505) { x *= 0.5; ++e; }
    const double s = (x - 1.0) / (x + 1.0), s2 = s * s;
    double p = 1.0 / 21.0;
    p = p * s2 + 1.0 / 19.0; p = p * s2 + 1.0 / 17.0; p = p * s2 + 1.0 / 15.0; p = p * s2 + 1.0 / 13.0;
    p = p * s2 + 1.0 / 11.0; p = p * s2 + 1.0 / 9.0; p = p * s2 + 1.0 / 7.0; p = p * s2 + 1.0 / 5.0;
    p = p * s2 + 1.0 / 3.0; p = p * s2 + 1.0;
    return 2.0 * s * p + (double) e * 0.69314718055994531;
}

// Standard normal deviate for noise sample n of a stream: Marsaglia's polar method on a splitmix64 sequence keyed
// by (seed, n), so that any sample can be drawn on its own (one device thread per sample).
M17G_FN double gauss_at(uint64_t noise_seed, uint64_t n)
{
    uint64_t z = noise_seed ^ (n * 0xD1342543DE82EF95ull);
    for (;;) {
        const double u = 2.0 * ((double) (splitmix64(z) >> 11) * (1.0 / 9007199254740992.0)) - 1.0;
        const double v = 2.0 * ((double) (splitmix64(z) >> 11) * (1.0 / 9007199254740992.0)) - 1.0;
        const double r = u * u + v * v;
        if (r >= 1.0 || r == 0.0) continue;
        return u * sqrt(-2.0 * log_pos(r) / r);
    }
}

// ---- FEC building blocks ---------------------------------------------------

// CRC-16, polynomial 0x5935, initial value 0xFFFF, MSB first, no reflection, no final xor
// (CRC16.h as instantiated in m17-mod.cpp:321; KAT "123456789" -> 0x772B).
M17G_FN uint16_t crc16_m17(const uint8_t* d, size_t n)
{
    uint16_t reg = 0xFFFF;
    for (size_t i = 0; i != n; ++i) {
        reg ^= uint16_t(d[i]) << 8;
        for (int b = 0; b != 8; ++b) reg = (reg & 0x8000) ? uint16_t((reg << 1) ^ 0x5935) : uint16_t(reg << 1);
    }
    return reg;
}

// CRC-16/X-25 (reflected 0x1021, init/xorout 0xFFFF): the packet FCS the reference CLI checks with residue
// 0x0f47 (apps/m17-demod.cpp:219-223).
M17G_FN uint16_t crc16_x25(const uint8_t* d, size_t n)
{
    uint16_t reg = 0xFFFF;
    for (size_t i = 0; i != n; ++i) {
        reg ^= d[i];
        for (int b = 0; b != 8; ++b) reg = (reg & 1) ? uint16_t((reg >> 1) ^ 0x8408) : uint16_t(reg >> 1);
    }
    return uint16_t(~reg);
}

// Golay(24,12), generator 0xC75 (Golay24.h:91,184-201): 11 check bits by polynomial division, then overall parity.
M17G_FN uint32_t golay24(uint16_t data)
{
    uint32_t cw = data & 0xFFF;
    for (int i = 0; i != 12; ++i) { if (cw & 1) cw ^= 0xC75; cw >>= 1; }
    uint32_t c23 = cw | (uint32_t(data & 0xFFF) << 11);
    return (c23 << 1) | (uint32_t) (popcount32(c23) & 1);
}

// K=5 rate-1/2 convolutional code, G1=031 G2=027 (octal), 4 flush bits (Convolution.h:12-21,
// m17-mod.cpp:357-375).  Returns 2*(nbits+4) coded bits.
M17G_FN size_t conv_encode(const uint8_t* bits, size_t nbits, uint8_t* out)
{
    uint32_t mem = 0;
    size_t k = 0;
    for (size_t i = 0; i != nbits + 4; ++i) {
        uint32_t x = i < nbits ? bits[i] : 0;
        mem = ((mem << 1) | x) & 0x1F;
        out[k++] = (uint8_t) (popcount32(mem & 031) & 1);
        out[k++] = (uint8_t) (popcount32(mem & 027) & 1);
    }
    return k;
}

// Puncture patterns: P1 (Trellis.h:17-30) drops every coded bit with index = 2 (mod 4) of a 61-bit period,
// P2 (:33-35) the 12th of every 12, P3 (:39-40) the 8th of every 8.  Keep coded bit i when the pattern allows;
// stop after `cap` outputs (Util.h:193-211).
M17G_FN size_t puncture(const uint8_t* in, size_t n, uint8_t* out, size_t cap, int which)
{
    const size_t plen = which == 1 ? 61 : which == 2 ? 12 : 8;
    size_t k = 0, p = 0;
    for (size_t i = 0; i != n && k != cap; ++i) {
        const bool keep = which == 1 ? (p % 4 != 2) : (p != plen - 1);
        if (keep) out[k++] = in[i];
        if (++p == plen) p = 0;
    }
    return k;
}

M17G_FN void bytes_to_bits(const uint8_t* bytes, size_t nbits, uint8_t* bits)
{
    for (size_t i = 0; i != nbits; ++i) bits[i] = (bytes[i >> 3] >> (7 - (i & 7))) & 1;
}

// QPP interleaver 45 i + 92 i^2 mod 368 (PolynomialInterleaver.h:21-34) then xor with the decorrelation
// sequence (M17Randomizer.h:51-57).
M17G_FN void interleave_randomize(const uint8_t* in /* 368 */, uint8_t* out /* 368 */)
{
    for (size_t i = 0; i != 368; ++i) out[(45 * i + 92 * i * i) % 368] = in[i];
    for (size_t i = 0; i != 368; ++i) out[i] ^= (dc_byte((int) (i >> 3)) >> (7 - (i & 7))) & 1;
}

M17G_FN void lsf_frame_bits(const uint8_t* lsf /* 30 */, uint8_t* out /* 368 */)
{
    uint8_t bits[240], coded[488], punct[368];
    bytes_to_bits(lsf, 240, bits);
    conv_encode(bits, 240, coded);
    puncture(coded, 488, punct, 368, 1);
    interleave_randomize(punct, out);
}

M17G_FN void lich_segment_bits(const uint8_t* lsf, int idx, uint8_t* out /* 96 */)
{
    const uint8_t* s = lsf + 5 * idx;
    uint16_t w[4] = {uint16_t((s[0] << 4) | (s[1] >> 4)), uint16_t(((s[1] & 0x0F) << 8) | s[2]),
                     uint16_t((s[3] << 4) | (s[4] >> 4)), uint16_t(((s[4] & 0x0F) << 8) | (idx << 5))};
    for (int k = 0; k != 4; ++k) {
        uint32_t cw = golay24(w[k]);
        for (int b = 0; b != 24; ++b) out[24 * k + b] = (cw >> (23 - b)) & 1;
    }
}

M17G_FN void stream_frame_bits(const uint8_t* lsf, int lich_idx, uint16_t fn, const uint8_t* payload /* 16 */, uint8_t* out)
{
    uint8_t data[18], bits[144], coded[296], frame[368];
    data[0] = uint8_t(fn >> 8); data[1] = uint8_t(fn);
    memcpy(data + 2, payload, 16);
    bytes_to_bits(data, 144, bits);
    conv_encode(bits, 144, coded);
    lich_segment_bits(lsf, lich_idx, frame);
    puncture(coded, 296, frame + 96, 272, 2);
    interleave_randomize(frame, out);
}

// 206 information bits: 25 data bytes + the top 6 bits of the control byte (EOF flag + 5-bit counter), as
// M17FrameDecoder.h:299-315 unpacks them.
M17G_FN void packet_frame_bits(const uint8_t* data26, uint8_t* out)
{
    uint8_t bits[208], coded[420], punct[368];
    bytes_to_bits(data26, 208, bits);
    conv_encode(bits, 206, coded);
    puncture(coded, 420, punct, 368, 3);
    interleave_randomize(punct, out);
}

M17G_FN int prbs9(uint16_t& st)   // Util.h:354-359
{
    int r = ((st >> 8) ^ (st >> 4)) & 1;
    st = uint16_t(((st << 1) | r) & 0x1FF);
    return r;
}

M17G_FN void bert_frame_bits(uint16_t& st, uint8_t* out)
{
    uint8_t bits[197], coded[402], punct[368];
    for (int i = 0; i != 197; ++i) bits[i] = uint8_t(prbs9(st));
    conv_encode(bits, 197, coded);
    puncture(coded, 402, punct, 368, 2);   // drops the 369th bit like Util.h:201
    interleave_randomize(punct, out);
}

M17G_FN void encode_callsign(const char* call, uint8_t* out /* 6 */)
{
    uint64_t enc = 0;
    size_t n = 0;
    while (call[n]) ++n;
    if (n > 9) n = 9;
    for (size_t i = n; i-- != 0;) {
        char c = call[i];
        enc *= 40;
        if (c >= 'A' && c <= 'Z') enc += uint64_t(c - 'A' + 1);
        else if (c >= '0' && c <= '9') enc += uint64_t(c - '0' + 27);
        else if (c == '-') enc += 37;
        else if (c == '/') enc += 38;
        else if (c == '.') enc += 39;
    }
    for (int i = 0; i != 6; ++i) out[i] = uint8_t(enc >> (8 * (5 - i)));
}

// ---- symbol stream assembly ------------------------------------------------

M17G_FN int8_t dibit_symbol(int b) { b &= 3; return (int8_t) (b == 0 ? 1 : b == 1 ? 3 : b == 2 ? -1 : -3); }

// Appends symbols to a caller-provided timeline; `n` keeps counting past `cap` (nothing is written there).
struct SymbolWriter {
    int8_t* s;
    size_t cap, n;
    M17G_FN void put(int8_t v) { if (n < cap) s[n] = v; ++n; }
    M17G_FN void bytes(const uint8_t* p, size_t count)
    {
        for (size_t i = 0; i != count; ++i)
            for (int k = 6; k >= 0; k -= 2) put(dibit_symbol(p[i] >> k));
    }
    M17G_FN void bits(const uint8_t* b, size_t count)
    {
        for (size_t i = 0; i != count; i += 2) put(dibit_symbol((b[i] << 1) | b[i + 1]));
    }
    M17G_FN void silence(size_t count) { for (size_t i = 0; i != count; ++i) put(0); }
};

M17G_FN void random_callsign(Rng& rng, char* out)
{
    int n = 4 + int(rng.below(3));
    for (int i = 0; i != n; ++i) {
        const uint32_t k = rng.below(i == 0 ? 26 : 36);
        out[i] = (char) (k < 26 ? 'A' + k : '0' + (k - 26));
    }
    out[n] = 0;
}

// One transmission appended to the timeline.
M17G_FN void transmission(Rng& rng, int kind, int n_frames, int encapsulated, SymbolWriter& w, m17gen_truth* truth)
{
    uint8_t frame[368];
    const uint8_t sync_lsf[2] = {0x55, 0xF7}, sync_stream[2] = {0xFF, 0x5D}, sync_packet[2] = {0x75, 0xFF},
                  sync_bert[2] = {0xDF, 0x55}, sync_eot[2] = {0x55, 0x5D};

    for (int i = 0; i != 48; ++i) { const uint8_t pre = 0x77; w.bytes(&pre, 1); }

    uint8_t lsf[30];
    memset(lsf, 0, 30);
    memset(lsf, 0xFF, 6);                          // broadcast destination
    char call[10];
    random_callsign(rng, call);
    encode_callsign(call, lsf + 6);
    int can = int(rng.below(16));
    if (kind == M17GEN_STREAM) {                   // m17-mod.cpp:338-340
        lsf[12] = uint8_t(can >> 1);
        lsf[13] = uint8_t(5 | ((can & 1) << 7));
    } else if (kind == M17GEN_BERT) {              // m17-mod.cpp:341-344
        lsf[12] = 0; lsf[13] = 1;
    } else {                                       // packet: type bit0 = 0, bits 2..1 = 1 RAW / 2 ENCAPSULATED
        lsf[12] = uint8_t(can >> 1);
        lsf[13] = uint8_t(((encapsulated ? 2 : 1) << 1) | ((can & 1) << 7));
    }
    for (int i = 14; i != 28; ++i) lsf[i] = uint8_t(rng.next());
    uint16_t crc = crc16_m17(lsf, 28);
    lsf[28] = uint8_t(crc >> 8); lsf[29] = uint8_t(crc);
    if (truth && truth->n_transmissions == 0) memcpy(truth->first_lsf, lsf, 30);
    if (truth) truth->n_transmissions += 1;

    lsf_frame_bits(lsf, frame);
    w.bytes(sync_lsf, 2);
    w.bits(frame, 368);

    if (kind == M17GEN_STREAM) {
        for (int f = 0; f != n_frames; ++f) {
            uint8_t payload[16];
            for (int i = 0; i != 16; ++i) payload[i] = uint8_t(rng.next());
            uint16_t fn = uint16_t(f & 0x7FFF);
            if (f == n_frames - 1) fn |= 0x8000;   // m17-mod.cpp:618-620
            stream_frame_bits(lsf, f % 6, fn, payload, frame);
            w.bytes(sync_stream, 2);
            w.bits(frame, 368);
            if (truth) truth->n_stream_frames += 1;
        }
    } else if (kind == M17GEN_PACKET) {
        // n_frames (<= 32) packet frames: 25*(n-1) + k bytes incl. the X.25 FCS.
        int last = 3 + int(rng.below(23));          // bytes in the final frame (>= 2 for the FCS)
        size_t total = size_t(25) * size_t(n_frames - 1) + size_t(last);
        uint8_t pkt[25 * 32 + 26];
        memset(pkt, 0, sizeof(pkt));
        for (size_t i = 0; i + 2 < total; ++i) pkt[i] = uint8_t(rng.next());
        uint16_t fcs = crc16_x25(pkt, total - 2);
        pkt[total - 2] = uint8_t(fcs & 0xFF); pkt[total - 1] = uint8_t(fcs >> 8);
        for (int f = 0; f != n_frames; ++f) {
            uint8_t d[26];
            memset(d, 0, 26);
            bool fin = f == n_frames - 1;
            size_t nb = fin ? size_t(last) : 25;
            memcpy(d, pkt + 25 * f, nb);
            d[25] = fin ? uint8_t(0x80 | (nb << 2)) : uint8_t((f & 0x1F) << 2);
            packet_frame_bits(d, frame);
            w.bytes(sync_packet, 2);
            w.bits(frame, 368);
            if (truth) truth->n_packet_frames += 1;
        }
    } else {
        uint16_t st = 1;
        for (int f = 0; f != n_frames; ++f) {
            bert_frame_bits(st, frame);
            w.bytes(sync_bert, 2);
            w.bits(frame, 368);
            if (truth) truth->n_bert_frames += 1;
        }
    }
    w.bytes(sync_eot, 2);      // m17-mod.cpp:297-305: 8 EOT symbols + 40 flush symbols
    w.silence(40);
}

// What the three stages of a stream share.
struct Plan {
    double rate, phase;        // TX samples per RX sample; initial fractional sample phase
    size_t tx_len, n_sym;      // TX-domain samples needed, symbols in the timeline
    double amp;
    uint64_t noise_seed;
};

M17G_FN Plan make_plan(const m17gen_params* p, size_t n_samples, Rng& rng)
{
    Plan q;
    q.rate = 1.0 + double(p->skew_ppm) * 1e-6;
    q.phase = p->phase >= 0 ? double(p->phase) : rng.uniform() * 10.0;
    q.tx_len = size_t(double(n_samples) * q.rate + q.phase) + 32;
    q.n_sym = q.tx_len / 10 + 2;
    q.amp = 7168.0 * double(p->gain) * (p->invert ? -1.0 : 1.0);
    q.noise_seed = p->seed ^ 0xA5A5A5A55A5A5A5Aull;
    return q;
}

// Upper bound on n_sym for buffer sizing (skew within +-5 %).
M17G_FN size_t max_symbols(size_t n_samples) { return size_t(double(n_samples) * 1.05 + 10.0 + 32.0) / 10 + 3; }

// Stage 1: the symbol timeline (+-1, +-3, 0 = no carrier), sym[0 .. n_sym).
M17G_FN void build_timeline(const m17gen_params* p, const Plan& q, Rng& rng, int8_t* sym, m17gen_truth* truth)
{
    SymbolWriter w{sym, q.n_sym, 0};
    const size_t n_sym = q.n_sym;
    w.silence(size_t(p->lead_in > 0 ? p->lead_in : 0) / 10);
    const size_t overhead = 192 + 192 + 48;     // preamble + LSF + EOT symbols
    const size_t tail = size_t(p->tail > 0 ? p->tail : 0) / 10;
    bool first = true;
    while (true) {
        size_t room = n_sym > w.n + tail ? n_sym - w.n - tail : 0;
        if (room < overhead + 192) break;
        int max_frames = int((room - overhead) / 192);
        int kind = p->mode;
        int nf = p->n_frames;
        if (p->mode == M17GEN_MIXED) {
            uint32_t r = rng.below(100);
            if (r < 50) { kind = M17GEN_STREAM; nf = 8 + int(rng.below(40)); }
            else if (r < 75) { kind = M17GEN_PACKET; nf = 1 + int(rng.below(32)); }
            else if (r < 90) { kind = M17GEN_BERT; nf = 4 + int(rng.below(24)); }
            else { kind = M17GEN_STREAM; nf = 1 + int(rng.below(5)); }
        }
        if (nf <= 0 || nf > max_frames) nf = max_frames;
        if (kind == M17GEN_PACKET && nf > 32) nf = 32;
        if (!first && !p->repeat && p->mode != M17GEN_MIXED) break;
        transmission(rng, kind, nf, p->packet_encapsulated, w, truth);
        first = false;
        if (!p->repeat && p->mode != M17GEN_MIXED) break;
        w.silence(size_t(p->gap > 0 ? p->gap : 0) / 10 + rng.below(192));
    }
    while (w.n < n_sym) w.put(0);
}

// TX-domain sample m: RRC pulse shaping at 10 samples/symbol (m17-mod.cpp:204-224), w[m] = sum_k sym[k] tap[m - 10 k].
M17G_FN double tx_sample(const int8_t* sym, size_t n_sym, double amp, size_t m)
{
    double acc = 0;
    const size_t k_hi = m / 10;
    for (size_t j = 0; j != 15 && j <= k_hi; ++j) {
        const size_t k = k_hi - j;
        if (k < n_sym && sym[k]) acc += (double) sym[k] * rrc_tap(int(m - 10 * k));
    }
    return acc * amp;
}

// Stage 2: RMS of the modulated signal over the symbol periods that carry one (sequential sum: one thread).
M17G_FN double signal_rms(const int8_t* sym, const Plan& q)
{
    double pw = 0;
    size_t pw_n = 0;
    for (size_t m = 0; m != q.tx_len + 8; ++m) {
        const size_t k_hi = m / 10;
        if (k_hi < q.n_sym && sym[k_hi]) {
            const double t = tx_sample(sym, q.n_sym, q.amp, m);
            pw += t * t; ++pw_n;
        }
    }
    return pw_n ? sqrt(pw / double(pw_n)) : 0.0;
}

// AWGN standard deviation from the signal's RMS; snr_lin = 10^(snr_db / 20), computed on the host.
M17G_FN double noise_sigma(const m17gen_params* p, double sig_rms, double snr_lin)
{
    double sigma = p->snr_db >= 200.f ? double(p->idle_noise_rms) : sig_rms / snr_lin;
    if (p->snr_db < 200.f && sigma < double(p->idle_noise_rms)) sigma = double(p->idle_noise_rms);
    return sigma;
}

// Stage 3: receiver sample n: resample to the receiver clock (6-point Lagrange), add DC + AWGN, quantise.
// tx(m) = TX-domain sample m (tx_sample(), or a table of its values).
template <typename Tx>
M17G_FN int16_t rx_sample(const m17gen_params* p, const Plan& q, double sigma, size_t n, Tx tx)
{
    const double t = double(n) * q.rate + q.phase + 2.0;    // +2: room for the left interpolation points
    const size_t i0 = size_t(t);
    const double f = t - double(i0);
    double v;
    if (p->skew_ppm == 0.f && f == 0.0) {
        v = tx(i0);
    } else {
        // nodes at -2..3 relative to i0
        const double x = f;
        const double c0 = -x * (x - 1) * (x + 1) * (x - 2) * (x - 3) / 120.0;
        const double c1 = x * (x + 2) * (x - 1) * (x - 2) * (x - 3) / 24.0;
        const double c2 = -(x + 2) * (x + 1) * (x - 1) * (x - 2) * (x - 3) / 12.0;
        const double c3 = x * (x + 2) * (x + 1) * (x - 2) * (x - 3) / 12.0;
        const double c4 = -x * (x + 2) * (x + 1) * (x - 1) * (x - 3) / 24.0;
        const double c5 = x * (x + 2) * (x + 1) * (x - 1) * (x - 2) / 120.0;
        v = c0 * tx(i0 - 2) + c1 * tx(i0 - 1) + c2 * tx(i0) + c3 * tx(i0 + 1) + c4 * tx(i0 + 2) + c5 * tx(i0 + 3);
    }
    v += double(p->dc_offset);
    if (sigma > 0) v += sigma * gauss_at(q.noise_seed, (uint64_t) n);
    long long r = llrint(v);
    r = r < -32768 ? -32768 : (r > 32767 ? 32767 : r);
    return (int16_t) r;
}

} // namespace m17gen
