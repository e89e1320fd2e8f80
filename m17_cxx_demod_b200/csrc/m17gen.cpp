// m17gen.cpp -- synthetic M17 baseband generator (see include/m17gen.h).
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
// Channel impairments (AWGN, clock skew, DC offset, gain) are this file's own.
// CPU only; not on the demodulation path.

#include "m17gen.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace {

// Root-raised-cosine taps (alpha 0.5, 10 samples/symbol, 150 taps incl. the
// trailing zero): the same design the reference ships (apps/m17-mod.cpp:36-38);
// only the first 75 are stored, the filter is symmetric about tap 74.
const double kRrcHalf[75] = {
    0.0029364388513841593, 0.0031468394550958484, 0.002699564567597445, 0.001661182944400927,
    0.00023319405581230247, -0.0012851320781224025, -0.0025577136087664687, -0.0032843366522956313,
    -0.0032697038088887226, -0.0024733964729590865, -0.0010285696910973807, 0.0007766690889758685,
    0.002553421969211845, 0.0038920145144327816, 0.004451886520053017, 0.00404219185231544,
    0.002674727068399207, 0.0005756567993179152, -0.0018493784971116507, -0.004092346891623224,
    -0.005648131453822014, -0.006126925416243605, -0.005349511529163396, -0.003403189203405097,
    -0.0006430502751187517, 0.002365929161655135, 0.004957956568090113, 0.006506845894531803,
    0.006569574194782443, 0.0050017573119839134, 0.002017321931508163, -0.0018256054303579805,
    -0.00571615173291049, -0.008746639552588416, -0.010105075751866371, -0.009265784007800534,
    -0.006136551625729697, -0.001125978562075172, 0.004891777252042491, 0.01071805138282269,
    0.01505751553351295, 0.01679337935001369, 0.015256245142156299, 0.01042830577908502,
    0.003031522725559901, -0.0055333532968188165, -0.013403099825723372, -0.018598682349642525,
    -0.01944761739590459, -0.015005271935951746, -0.0053887880354343935, 0.008056525910253532,
    0.022816244158307273, 0.035513467692208076, 0.04244131815783876, 0.04025481153629372,
    0.02671818654865632, 0.0013810216516704976, -0.03394615682795165, -0.07502635967975885,
    -0.11540977897637611, -0.14703962203941534, -0.16119995609538576, -0.14969512896336504,
    -0.10610329539459686, -0.026921412469634916, 0.08757875030779196, 0.23293327870303457,
    0.4006012210123992, 0.5786324696325503, 0.7528286479934068, 0.908262741447522,
    1.0309661131633199, 1.1095611856548013, 1.1366197723675815};

inline double rrc_tap(int i)   // i in [0,150)
{
    if (i < 0 || i >= 149) return 0.0;
    return i <= 74 ? kRrcHalf[i] : kRrcHalf[148 - i];
}

// M17 decorrelation sequence (M17Randomizer.h:16-22; M17 spec table).
const uint8_t kDC[46] = {
    0xd6, 0xb5, 0xe2, 0x30, 0x82, 0xFF, 0x84, 0x62, 0xba, 0x4e, 0x96, 0x90, 0xd8, 0x98, 0xdd, 0x5d,
    0x0c, 0xc8, 0x52, 0x43, 0x91, 0x1d, 0xf8, 0x6e, 0x68, 0x2F, 0x35, 0xda, 0x14, 0xea, 0xcd, 0x76,
    0x19, 0x8d, 0xd5, 0x80, 0xd1, 0x33, 0x87, 0x13, 0x57, 0x18, 0x2d, 0x29, 0x78, 0xc3};

struct Rng {
    uint64_t s[4];
    bool have_spare = false;
    double spare = 0;
    explicit Rng(uint64_t seed)
    {
        uint64_t z = seed;
        for (auto& x : s) {   // splitmix64
            z += 0x9E3779B97F4A7C15ull;
            uint64_t y = z;
            y = (y ^ (y >> 30)) * 0xBF58476D1CE4E5B9ull;
            y = (y ^ (y >> 27)) * 0x94D049BB133111EBull;
            x = y ^ (y >> 31);
        }
    }
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t next()   // xoshiro256**
    {
        uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    double uniform() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
    uint32_t below(uint32_t n) { return uint32_t((next() >> 32) * uint64_t(n) >> 32); }
    double gauss()
    {
        if (have_spare) { have_spare = false; return spare; }
        double u, v, r;
        do { u = 2 * uniform() - 1; v = 2 * uniform() - 1; r = u * u + v * v; } while (r >= 1.0 || r == 0.0);
        double f = std::sqrt(-2.0 * std::log(r) / r);
        spare = v * f; have_spare = true;
        return u * f;
    }
};

// ---- FEC building blocks ---------------------------------------------------

// CRC-16, polynomial 0x5935, initial value 0xFFFF, MSB first, no reflection,
// no final xor (CRC16.h as instantiated in m17-mod.cpp:321; KAT "123456789" -> 0x772B).
uint16_t crc16_m17(const uint8_t* d, size_t n)
{
    uint16_t reg = 0xFFFF;
    for (size_t i = 0; i != n; ++i) {
        reg ^= uint16_t(d[i]) << 8;
        for (int b = 0; b != 8; ++b) reg = (reg & 0x8000) ? uint16_t((reg << 1) ^ 0x5935) : uint16_t(reg << 1);
    }
    return reg;
}

// CRC-16/X-25 (reflected 0x1021, init/xorout 0xFFFF): the packet FCS the
// reference CLI checks with residue 0x0f47 (apps/m17-demod.cpp:219-223).
uint16_t crc16_x25(const uint8_t* d, size_t n)
{
    uint16_t reg = 0xFFFF;
    for (size_t i = 0; i != n; ++i) {
        reg ^= d[i];
        for (int b = 0; b != 8; ++b) reg = (reg & 1) ? uint16_t((reg >> 1) ^ 0x8408) : uint16_t(reg >> 1);
    }
    return uint16_t(~reg);
}

// Golay(24,12), generator 0xC75 (Golay24.h:91,184-201): 11 check bits by
// polynomial division, then overall parity.
uint32_t golay24(uint16_t data)
{
    uint32_t cw = data & 0xFFF;
    for (int i = 0; i != 12; ++i) { if (cw & 1) cw ^= 0xC75; cw >>= 1; }
    uint32_t c23 = cw | (uint32_t(data & 0xFFF) << 11);
    return (c23 << 1) | (__builtin_popcount(c23) & 1);
}

// K=5 rate-1/2 convolutional code, G1=031 G2=027 (octal), 4 flush bits
// (Convolution.h:12-21, m17-mod.cpp:357-375).  Returns 2*(nbits+4) coded bits.
size_t conv_encode(const uint8_t* bits, size_t nbits, uint8_t* out)
{
    uint32_t mem = 0;
    size_t k = 0;
    for (size_t i = 0; i != nbits + 4; ++i) {
        uint32_t x = i < nbits ? bits[i] : 0;
        mem = ((mem << 1) | x) & 0x1F;
        out[k++] = __builtin_popcount(mem & 031) & 1;
        out[k++] = __builtin_popcount(mem & 027) & 1;
    }
    return k;
}

// Keep coded bit i when pattern[i mod plen] != 0; stop after `cap` outputs
// (Util.h:193-211).
size_t puncture(const uint8_t* in, size_t n, uint8_t* out, size_t cap, const uint8_t* pat, size_t plen)
{
    size_t k = 0, p = 0;
    for (size_t i = 0; i != n && k != cap; ++i) {
        if (pat[p]) out[k++] = in[i];
        if (++p == plen) p = 0;
    }
    return k;
}

struct Patterns {
    uint8_t p1[61], p2[12], p3[8];
    Patterns()
    {
        for (int i = 0; i != 61; ++i) p1[i] = (i % 4 == 2) ? 0 : 1;   // Trellis.h:17-30
        for (int i = 0; i != 12; ++i) p2[i] = i != 11;                // Trellis.h:33-35
        for (int i = 0; i != 8; ++i) p3[i] = i != 7;                  // Trellis.h:39-40
    }
};
const Patterns kPat;

void bytes_to_bits(const uint8_t* bytes, size_t nbits, uint8_t* bits)
{
    for (size_t i = 0; i != nbits; ++i) bits[i] = (bytes[i >> 3] >> (7 - (i & 7))) & 1;
}

// QPP interleaver 45 i + 92 i^2 mod 368 (PolynomialInterleaver.h:21-34) then
// xor with the decorrelation sequence (M17Randomizer.h:51-57).
void interleave_randomize(const uint8_t in[368], uint8_t out[368])
{
    for (size_t i = 0; i != 368; ++i) out[(45 * i + 92 * i * i) % 368] = in[i];
    for (size_t i = 0; i != 368; ++i) out[i] ^= (kDC[i >> 3] >> (7 - (i & 7))) & 1;
}

void lsf_frame_bits(const uint8_t lsf[30], uint8_t out[368])
{
    uint8_t bits[240], coded[488], punct[368];
    bytes_to_bits(lsf, 240, bits);
    conv_encode(bits, 240, coded);
    puncture(coded, 488, punct, 368, kPat.p1, 61);
    interleave_randomize(punct, out);
}

void lich_segment_bits(const uint8_t lsf[30], int idx, uint8_t out[96])
{
    const uint8_t* s = lsf + 5 * idx;
    uint16_t w[4] = {uint16_t((s[0] << 4) | (s[1] >> 4)), uint16_t(((s[1] & 0x0F) << 8) | s[2]),
                     uint16_t((s[3] << 4) | (s[4] >> 4)), uint16_t(((s[4] & 0x0F) << 8) | (idx << 5))};
    for (int k = 0; k != 4; ++k) {
        uint32_t cw = golay24(w[k]);
        for (int b = 0; b != 24; ++b) out[24 * k + b] = (cw >> (23 - b)) & 1;
    }
}

void stream_frame_bits(const uint8_t lsf[30], int lich_idx, uint16_t fn, const uint8_t payload[16], uint8_t out[368])
{
    uint8_t data[18], bits[144], coded[296], frame[368];
    data[0] = uint8_t(fn >> 8); data[1] = uint8_t(fn);
    std::memcpy(data + 2, payload, 16);
    bytes_to_bits(data, 144, bits);
    conv_encode(bits, 144, coded);
    lich_segment_bits(lsf, lich_idx, frame);
    puncture(coded, 296, frame + 96, 272, kPat.p2, 12);
    interleave_randomize(frame, out);
}

// 206 information bits: 25 data bytes + the top 6 bits of the control byte
// (EOF flag + 5-bit counter), as M17FrameDecoder.h:299-315 unpacks them.
void packet_frame_bits(const uint8_t data26[26], uint8_t out[368])
{
    uint8_t bits[208], coded[420], punct[368];
    bytes_to_bits(data26, 208, bits);
    conv_encode(bits, 206, coded);
    puncture(coded, 420, punct, 368, kPat.p3, 8);
    interleave_randomize(punct, out);
}

inline int prbs9(uint16_t& st)   // Util.h:354-359
{
    int r = ((st >> 8) ^ (st >> 4)) & 1;
    st = uint16_t(((st << 1) | r) & 0x1FF);
    return r;
}

void bert_frame_bits(uint16_t& st, uint8_t out[368])
{
    uint8_t bits[197], coded[402], punct[368];
    for (auto& b : bits) b = uint8_t(prbs9(st));
    conv_encode(bits, 197, coded);
    puncture(coded, 402, punct, 368, kPat.p2, 12);   // drops the 369th bit like Util.h:201
    interleave_randomize(punct, out);
}

void encode_callsign(const char* call, uint8_t out[6])
{
    uint64_t enc = 0;
    size_t n = std::strlen(call);
    if (n > 9) n = 9;
    for (size_t i = n; i-- != 0;) {
        char c = call[i];
        enc *= 40;
        if (c >= 'A' && c <= 'Z') enc += c - 'A' + 1;
        else if (c >= '0' && c <= '9') enc += c - '0' + 27;
        else if (c == '-') enc += 37;
        else if (c == '/') enc += 38;
        else if (c == '.') enc += 39;
    }
    for (int i = 0; i != 6; ++i) out[i] = uint8_t(enc >> (8 * (5 - i)));
}

// ---- symbol stream assembly ------------------------------------------------

inline int8_t dibit_symbol(int b) { static const int8_t m[4] = {1, 3, -1, -3}; return m[b & 3]; }

struct SymbolWriter {
    std::vector<int8_t>& s;
    void bytes(const uint8_t* p, size_t n)
    {
        for (size_t i = 0; i != n; ++i)
            for (int k = 6; k >= 0; k -= 2) s.push_back(dibit_symbol(p[i] >> k));
    }
    void bits(const uint8_t* b, size_t n)
    {
        for (size_t i = 0; i != n; i += 2) s.push_back(dibit_symbol((b[i] << 1) | b[i + 1]));
    }
    void silence(size_t n) { s.insert(s.end(), n, 0); }
};

const uint8_t kSyncLsf[2] = {0x55, 0xF7}, kSyncStream[2] = {0xFF, 0x5D}, kSyncPacket[2] = {0x75, 0xFF},
              kSyncBert[2] = {0xDF, 0x55}, kSyncEot[2] = {0x55, 0x5D};

void random_callsign(Rng& rng, char* out)
{
    static const char cs[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789";
    int n = 4 + int(rng.below(3));
    for (int i = 0; i != n; ++i) out[i] = cs[rng.below(i == 0 ? 26 : 36)];
    out[n] = 0;
}

// One transmission appended to `sym`; returns the LSF used.
void transmission(Rng& rng, int kind, int n_frames, int encapsulated, std::vector<int8_t>& sym, m17gen_truth* truth)
{
    SymbolWriter w{sym};
    uint8_t frame[368];

    std::array<uint8_t, 48> pre; pre.fill(0x77);
    w.bytes(pre.data(), 48);

    uint8_t lsf[30] = {0};
    std::memset(lsf, 0xFF, 6);                     // broadcast destination
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
    if (truth && truth->n_transmissions == 0) std::memcpy(truth->first_lsf, lsf, 30);
    if (truth) truth->n_transmissions += 1;

    lsf_frame_bits(lsf, frame);
    w.bytes(kSyncLsf, 2);
    w.bits(frame, 368);

    if (kind == M17GEN_STREAM) {
        for (int f = 0; f != n_frames; ++f) {
            uint8_t payload[16];
            for (auto& b : payload) b = uint8_t(rng.next());
            uint16_t fn = uint16_t(f & 0x7FFF);
            if (f == n_frames - 1) fn |= 0x8000;   // m17-mod.cpp:618-620
            stream_frame_bits(lsf, f % 6, fn, payload, frame);
            w.bytes(kSyncStream, 2);
            w.bits(frame, 368);
            if (truth) truth->n_stream_frames += 1;
        }
    } else if (kind == M17GEN_PACKET) {
        // n_frames packet frames: 25*(n-1) + k bytes incl. the X.25 FCS.
        int last = 3 + int(rng.below(23));          // bytes in the final frame (>= 2 for the FCS)
        size_t total = size_t(25) * (n_frames - 1) + last;
        std::vector<uint8_t> pkt(total);
        for (size_t i = 0; i + 2 < total; ++i) pkt[i] = uint8_t(rng.next());
        uint16_t fcs = crc16_x25(pkt.data(), total - 2);
        pkt[total - 2] = uint8_t(fcs & 0xFF); pkt[total - 1] = uint8_t(fcs >> 8);
        for (int f = 0; f != n_frames; ++f) {
            uint8_t d[26] = {0};
            bool fin = f == n_frames - 1;
            size_t nb = fin ? size_t(last) : 25;
            std::memcpy(d, pkt.data() + 25 * f, nb);
            d[25] = fin ? uint8_t(0x80 | (nb << 2)) : uint8_t((f & 0x1F) << 2);
            packet_frame_bits(d, frame);
            w.bytes(kSyncPacket, 2);
            w.bits(frame, 368);
            if (truth) truth->n_packet_frames += 1;
        }
    } else {
        uint16_t st = 1;
        for (int f = 0; f != n_frames; ++f) {
            bert_frame_bits(st, frame);
            w.bytes(kSyncBert, 2);
            w.bits(frame, 368);
            if (truth) truth->n_bert_frames += 1;
        }
    }
    w.bytes(kSyncEot, 2);      // m17-mod.cpp:297-305: 8 EOT symbols + 40 flush symbols
    w.silence(40);
}

} // namespace

extern "C" {

void m17gen_default_params(m17gen_params* p)
{
    std::memset(p, 0, sizeof(*p));
    p->seed = 0x4D3137;
    p->mode = M17GEN_STREAM;
    p->n_frames = 0;
    p->lead_in = 9600;
    p->tail = 4800;
    p->gap = 9600;
    p->snr_db = 40.f;
    p->gain = 1.f;
    p->phase = -1.f;
    p->idle_noise_rms = 8.f;
}

int m17gen_stream(const m17gen_params* p, int16_t* out, size_t n_samples, m17gen_truth* truth)
{
    if (!p || !out) return -1;
    if (truth) std::memset(truth, 0, sizeof(*truth));
    Rng rng(p->seed);
    Rng noise(p->seed ^ 0xA5A5A5A55A5A5A5Aull);

    const double rate = 1.0 + double(p->skew_ppm) * 1e-6;     // TX samples per RX sample
    const double phase = p->phase >= 0 ? double(p->phase) : rng.uniform() * 10.0;
    const size_t tx_len = size_t(double(n_samples) * rate + phase) + 32;   // TX-domain samples needed
    const size_t n_sym = tx_len / 10 + 2;

    // ---- symbol timeline -----------------------------------------------------
    std::vector<int8_t> sym;
    sym.reserve(n_sym + 4096);
    SymbolWriter w{sym};
    w.silence(size_t(std::max(0, p->lead_in)) / 10);
    const size_t overhead = 192 + 192 + 48;     // preamble + LSF + EOT symbols
    bool first = true;
    while (true) {
        size_t room = n_sym > sym.size() + size_t(std::max(0, p->tail)) / 10 ? n_sym - sym.size() - size_t(std::max(0, p->tail)) / 10 : 0;
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
        transmission(rng, kind, nf, p->packet_encapsulated, sym, truth);
        first = false;
        if (!p->repeat && p->mode != M17GEN_MIXED) break;
        w.silence(size_t(std::max(0, p->gap)) / 10 + rng.below(192));
    }
    if (sym.size() < n_sym) sym.resize(n_sym, 0);

    // ---- RRC pulse shaping at 10 samples/symbol (m17-mod.cpp:204-224) -----------
    const double amp = 7168.0 * double(p->gain) * (p->invert ? -1.0 : 1.0);
    std::vector<double> tx(tx_len + 8, 0.0);
    double pw = 0; size_t pw_n = 0;
    for (size_t m = 0; m != tx.size(); ++m) {
        // w[m] = sum_k sym[k] * tap[m - 10 k]
        double acc = 0;
        size_t k_hi = m / 10;
        for (size_t j = 0; j != 15 && j <= k_hi; ++j) {
            size_t k = k_hi - j;
            if (k < sym.size() && sym[k]) acc += sym[k] * rrc_tap(int(m - 10 * k));
        }
        tx[m] = acc * amp;
        if (k_hi < sym.size() && sym[k_hi]) { pw += tx[m] * tx[m]; ++pw_n; }
    }
    const double sig_rms = pw_n ? std::sqrt(pw / double(pw_n)) : 0.0;
    double sigma = p->snr_db >= 200.f ? double(p->idle_noise_rms)
                                      : sig_rms / std::pow(10.0, double(p->snr_db) / 20.0);
    if (p->snr_db < 200.f && sigma < double(p->idle_noise_rms)) sigma = double(p->idle_noise_rms);

    // ---- resample to the receiver clock (6-point Lagrange), add DC + AWGN --------
    for (size_t n = 0; n != n_samples; ++n) {
        double t = double(n) * rate + phase + 2.0;    // +2: room for the left interpolation points
        size_t i0 = size_t(t);
        double f = t - double(i0);
        double v;
        if (p->skew_ppm == 0.f && f == 0.0) {
            v = tx[i0];
        } else {
            // nodes at -2..3 relative to i0
            double x = f;
            double c0 = -x * (x - 1) * (x + 1) * (x - 2) * (x - 3) / 120.0;
            double c1 = x * (x + 2) * (x - 1) * (x - 2) * (x - 3) / 24.0;
            double c2 = -(x + 2) * (x + 1) * (x - 1) * (x - 2) * (x - 3) / 12.0;
            double c3 = x * (x + 2) * (x + 1) * (x - 2) * (x - 3) / 12.0;
            double c4 = -x * (x + 2) * (x + 1) * (x - 1) * (x - 3) / 24.0;
            double c5 = x * (x + 2) * (x + 1) * (x - 1) * (x - 2) / 120.0;
            v = c0 * tx[i0 - 2] + c1 * tx[i0 - 1] + c2 * tx[i0] + c3 * tx[i0 + 1] + c4 * tx[i0 + 2] + c5 * tx[i0 + 3];
        }
        v += double(p->dc_offset);
        if (sigma > 0) v += sigma * noise.gauss();
        long q = std::lrint(v);
        out[n] = int16_t(std::clamp<long>(q, -32768, 32767));
    }
    return 0;
}

int m17gen_batch(const m17gen_params* params, size_t n_streams, int16_t* out, size_t n_samples,
                 size_t stride, m17gen_truth* truth, int n_threads)
{
    if (n_threads < 1) n_threads = 1;
    std::vector<std::thread> pool;
    std::vector<int> rcs(size_t(n_threads), 0);
    for (int t = 0; t != n_threads; ++t) {
        pool.emplace_back([&, t] {
            for (size_t s = size_t(t); s < n_streams; s += size_t(n_threads)) {
                int rc = m17gen_stream(&params[s], out + s * stride, n_samples, truth ? &truth[s] : nullptr);
                if (rc) rcs[size_t(t)] = rc;
            }
        });
    }
    for (auto& th : pool) th.join();
    for (int rc : rcs) if (rc) return rc;
    return 0;
}

uint16_t m17gen_crc16(const uint8_t* data, size_t n) { return crc16_m17(data, n); }
uint32_t m17gen_golay24_encode(uint16_t data12) { return golay24(data12); }
void m17gen_lsf_frame_bits(const uint8_t lsf[30], uint8_t out368[368]) { lsf_frame_bits(lsf, out368); }
void m17gen_stream_frame_bits(const uint8_t lsf[30], int lich_index, uint16_t fn, const uint8_t payload16[16],
                              uint8_t out368[368]) { stream_frame_bits(lsf, lich_index, fn, payload16, out368); }
void m17gen_packet_frame_bits(const uint8_t data26[26], uint8_t out368[368]) { packet_frame_bits(data26, out368); }
void m17gen_bert_frame_bits(uint16_t* prbs_state, uint8_t out368[368]) { bert_frame_bits(*prbs_state, out368); }
void m17gen_encode_callsign(const char* call, uint8_t out6[6]) { encode_callsign(call, out6); }

} // extern "C"
