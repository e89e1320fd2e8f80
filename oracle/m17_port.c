/*
 * m17_port.c -- plain-C restatement of the reference demodulation path.
 * TEST INFRASTRUCTURE ONLY; see m17_port.h / m17_oracle.h for the contract.
 *
 * Every function cites the reference lines it follows (paths relative to
 * /root/reference/include/m17cxx/ unless stated).  Arithmetic is written out
 * operation by operation in the precision the reference's compiled code uses
 * (x86-64 baseline: IEEE float/double, no FMA contraction); build with
 * -ffp-contract=off (oracle/Makefile).
 */
#include "m17_port.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------- */
/* constants                                                                 */
/* ------------------------------------------------------------------------- */

/* M17Demodulator.h:79-118 detail::Taps<float>::rrc_taps (double literals narrowed
 * to float); symmetric about tap 74, tap 149 is 0. */
static const double kRrcHalf[75] = {
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

static float g_taps[150];

/* SlidingDFT.h:85-95: coeff = std::exp(-j * 2pi * f/48000) evaluated at run time in
 * complex<float> for f = 2400, 3600.  The four floats below are the values the
 * reference build produces on glibc 2.39 (read back through
 * m17ref_sdft_coeffs; tests/test_oracle_vs_ref.py re-checks them). */
static const uint32_t kSdftCoeffBits[4] = {0x3f737871u, 0xbe9e377au, 0x3f641901u, 0xbee87173u};
static float g_sdft_c[4];

/* Correlator.h:38-39 */
static const float kIirB[3] = {4.24433681e-05f, 8.48867363e-05f, 4.24433681e-05f};
static const float kIirA[3] = {1.0f, -1.98148851f, 0.98165828f};

/* M17Randomizer.h:16-22 */
static const uint8_t kDC[46] = {
    0xd6, 0xb5, 0xe2, 0x30, 0x82, 0xFF, 0x84, 0x62, 0xba, 0x4e, 0x96, 0x90, 0xd8, 0x98, 0xdd, 0x5d,
    0x0c, 0xc8, 0x52, 0x43, 0x91, 0x1d, 0xf8, 0x6e, 0x68, 0x2F, 0x35, 0xda, 0x14, 0xea, 0xcd, 0x76,
    0x19, 0x8d, 0xd5, 0x80, 0xd1, 0x33, 0x87, 0x13, 0x57, 0x18, 0x2d, 0x29, 0x78, 0xc3};

enum { ST_UNLOCKED, ST_LSF_SYNC, ST_STREAM_SYNC, ST_PACKET_SYNC, ST_BERT_SYNC, ST_SYNC_WAIT, ST_FRAME };
enum { DEC_LSF, DEC_STREAM, DEC_BASIC_PACKET, DEC_FULL_PACKET, DEC_BERT };
enum { SW_LSF, SW_STREAM, SW_PACKET, SW_BERT };
enum { RES_FAIL, RES_OK, RES_EOS, RES_INCOMPLETE, RES_PACKET_INCOMPLETE };

static float    g_llr_thr[43];
static int8_t   g_llr_val[43][2];
static int8_t   g_dc[368];
static uint16_t g_il[368];
static int8_t   g_p1[61], g_p2[12], g_p3[8];
static int16_t  g_vcost[8][2];
static uint32_t g_golay_corr[2048];
static uint16_t g_crc_init;
static int      g_ready = 0;

/* ------------------------------------------------------------------------- */
/* table construction                                                        */
/* ------------------------------------------------------------------------- */

/* Util.h:63-104 make_llr_map<float,4>: thresholds by repeated float addition of
 * inc = 1/7 starting from -3 + inc; comparisons `k + 1.0 < 0`, `k - 1.0 < 0` in double. */
static void make_llr_map(void)
{
    const int8_t limit = 7;
    const float inc = (float) (1.0 / (double) (float) limit);
    int8_t i = limit, j = limit;
    float k = (float) (-3.0 + (double) inc);
    for (int index = 0; index != 43; ++index) {
        g_llr_thr[index] = k;
        g_llr_val[index][0] = i;
        g_llr_val[index][1] = j;
        if ((double) k + 1.0 < 0) {
            j--;
            if (j == 0) j = -1;
            if (j < -limit) j = -limit;
        } else if ((double) k - 1.0 < 0) {
            i--;
            if (i == 0) i = -1;
            if (i < -limit) i = -limit;
        } else {
            j++;
            if (j == 0) j = 1;
            if (j > limit) j = limit;
        }
        k = k + inc;
    }
}

/* Golay24.h:102-112 */
static uint32_t golay_syndrome(uint32_t codeword)
{
    codeword &= 0xffffffu;
    for (int i = 0; i != 12; ++i) {
        if (codeword & 1) codeword ^= 0xC75;
        codeword >>= 1;
    }
    return codeword << 12;
}

/* Golay24.h:131-177: the reference sorts (syndrome, error pattern) pairs for all
 * patterns of weight <= 3 over 23 bits and binary-searches them.  The code is
 * perfect (1+23+253+1771 = 2048 = 2^11 distinct syndromes), so a direct table
 * indexed by the 11-bit syndrome returns the same entry lower_bound finds. */
static void make_golay_lut(void)
{
    memset(g_golay_corr, 0xFF, sizeof(g_golay_corr));
    g_golay_corr[golay_syndrome(0) >> 12] = 0;
    for (int i = 0; i != 23; ++i) {
        uint32_t v = 1u << i;
        g_golay_corr[golay_syndrome(v) >> 12] = v;
    }
    for (int i = 0; i != 22; ++i)
        for (int j = i + 1; j != 23; ++j) {
            uint32_t v = (1u << i) | (1u << j);
            g_golay_corr[golay_syndrome(v) >> 12] = v;
        }
    for (int i = 0; i != 21; ++i)
        for (int j = i + 1; j != 22; ++j)
            for (int k = j + 1; k != 23; ++k) {
                uint32_t v = (1u << i) | (1u << j) | (1u << k);
                g_golay_corr[golay_syndrome(v) >> 12] = v;
            }
}

static void m17p_setup(void)
{
    if (g_ready) return;
    for (int i = 0; i != 150; ++i)
        g_taps[i] = i == 149 ? 0.0f : (float) (i <= 74 ? kRrcHalf[i] : kRrcHalf[148 - i]);
    memcpy(g_sdft_c, kSdftCoeffBits, sizeof(g_sdft_c));
    make_llr_map();
    /* M17Randomizer.h:29-39 */
    for (int i = 0; i != 368; ++i) g_dc[i] = ((kDC[i >> 3] >> (7 - (i & 7))) & 1) ? -1 : 1;
    /* PolynomialInterleaver.h:21-24 */
    for (unsigned i = 0; i != 368; ++i) g_il[i] = (uint16_t) ((45u * i + 92u * i * i) % 368u);
    /* Trellis.h:17-40 */
    for (int i = 0, j = 2; i != 61; ++i) { if (i == j) { g_p1[i] = 0; j += 4; } else g_p1[i] = 1; }
    for (int i = 0; i != 12; ++i) g_p2[i] = i != 11;
    for (int i = 0; i != 8; ++i) g_p3[i] = i != 7;
    /* Viterbi.h:72-88 makeCost<Trellis<4,2>,4>: cost[i][j] = +7 if the coded bit
     * convolve_bit(poly_j, i << 1) is 1 else -7 (only butterflies i < 8 are used). */
    for (unsigned i = 0; i != 8; ++i) {
        g_vcost[i][0] = (__builtin_popcount(031u & (i << 1)) & 1) ? 7 : -7;
        g_vcost[i][1] = (__builtin_popcount(027u & (i << 1)) & 1) ? 7 : -7;
    }
    make_golay_lut();
    /* CRC16.h:21-34 reset() pre-conditioning of Init = 0xFFFF */
    {
        uint16_t reg = 0xFFFF;
        for (int i = 0; i != 16; ++i) {
            uint16_t bit = reg & 1;
            if (bit) reg ^= 0x5935;
            reg >>= 1;
            if (bit) reg |= 0x8000;
        }
        g_crc_init = reg;
    }
    g_ready = 1;
}

/* ------------------------------------------------------------------------- */
/* integer blocks                                                            */
/* ------------------------------------------------------------------------- */

/* CRC16.h:41-62 */
static uint16_t crc16_byte(uint8_t byte, uint16_t reg)
{
    for (int i = 0; i != 8; ++i) {
        uint16_t msb = reg & 0x8000;
        reg = (uint16_t) (((reg << 1) & 0xFFFF) | ((byte >> (7 - i)) & 1));
        if (msb) reg ^= 0x5935;
    }
    return reg;
}

static uint16_t crc16_get(uint16_t reg)
{
    for (int i = 0; i != 16; ++i) {
        uint16_t msb = reg & 0x8000;
        reg = (uint16_t) ((reg << 1) & 0xFFFF);
        if (msb) reg ^= 0x5935;
    }
    return reg;
}

static uint16_t crc16(const uint8_t* d, size_t n)
{
    uint16_t reg = g_crc_init;
    for (size_t i = 0; i != n; ++i) reg = crc16_byte(d[i], reg);
    return crc16_get(reg);
}

/* Golay24.h:184-201 */
static uint32_t golay_encode24(uint16_t data)
{
    uint32_t codeword = data;
    for (int i = 0; i != 12; ++i) {
        if (codeword & 1) codeword ^= 0xC75;
        codeword >>= 1;
    }
    codeword |= ((uint32_t) data << 11);
    return (codeword << 1) | (__builtin_popcount(codeword) & 1);
}

/* Golay24.h:203-222 */
static int golay_decode(uint32_t input, uint32_t* output)
{
    uint32_t syndrm = golay_syndrome(input >> 1);
    uint32_t bits = g_golay_corr[syndrm >> 12];
    uint32_t correction = bits << 1;
    *output = input ^ correction;
    return __builtin_popcount(syndrm) < 3 || !(__builtin_popcount(*output) & 1);
}

/* Util.h:169-190 */
static void depuncture(const int8_t* in, size_t n_in, int8_t* out, size_t n_out, const int8_t* p, size_t plen)
{
    size_t index = 0, pindex = 0;
    for (size_t i = 0; i != n_out && index < n_in; ++i) {
        if (!p[pindex++]) out[i] = 0;
        else out[i] = in[index++];
        if (pindex == plen) pindex = 0;
    }
}

/* Viterbi.h:128-236, Viterbi<Trellis<4,2>,4>::decode<IN,OUT>.  n_in soft bits
 * (0 = erasure), n_out decoded bits; returns round(min_metric / 7.f). */
static uint64_t viterbi_decode(const int8_t* in, size_t n_in, uint8_t* out, size_t n_out)
{
    int32_t prev[16], curr[16];
    uint16_t history[244];
    const int32_t MAX_METRIC = 0x7fffffff / 2;
    for (int i = 0; i != 16; ++i) prev[i] = MAX_METRIC;
    prev[0] = 0;
    size_t steps = n_in / 2;
    for (size_t t = 0; t != steps; ++t) {
        int16_t s0 = in[2 * t], s1 = in[2 * t + 1];
        int16_t cost0[8], cost1[8];
        for (int j = 0; j != 8; ++j) {
            cost0[j] = 0; cost1[j] = 0;
            if (s0) {
                cost0[j] = (int16_t) abs(g_vcost[j][0] - s0);
                cost1[j] = (int16_t) abs(g_vcost[j][0] + s0);
            }
            if (s1) {
                cost0[j] += (int16_t) abs(g_vcost[j][1] - s1);
                cost1[j] += (int16_t) abs(g_vcost[j][1] + s1);
            }
        }
        uint16_t hist = 0;
        for (int j = 0; j != 8; ++j) {
            int i0 = (2 * j) & 15, i1 = (2 * j + 1) & 15;     /* nextState_[j][0/1], Viterbi.h:25-37 */
            int32_t p0 = prev[j], p1 = prev[j + 8];
            int32_t m0 = p0 + cost0[j], m1 = p0 + cost1[j], m2 = p1 + cost1[j], m3 = p1 + cost0[j];
            int d0 = m0 > m2, d1 = m1 > m3;
            hist |= (uint16_t) (d0 << i0);
            hist |= (uint16_t) (d1 << i1);
            curr[i0] = d0 ? m2 : m0;
            curr[i1] = d1 ? m3 : m1;
        }
        history[t] = hist;
        memcpy(prev, curr, sizeof(prev));
    }
    size_t min_element = 0;
    int32_t min_cost = prev[0];
    for (size_t i = 0; i != 16; ++i)
        if (prev[i] < min_cost) { min_cost = prev[i]; min_element = i; }
    uint64_t cost = (uint64_t) roundf((float) min_cost / 7.0f);

    /* chainback: prevState_[s][v] = (s >> 1) | (v << 3)  (Viterbi.h:47-64) */
    size_t next = min_element, index = steps, o = n_out;
    for (size_t t = steps; t-- != 0 && o != 0;) {
        int v = (history[t] >> next) & 1;
        if (index-- <= n_out) out[--o] = (uint8_t) (next & 1);
        next = (next >> 1) | ((size_t) v << 3);
    }
    return cost;
}

/* Util.h:300-318 (second overload) */
static void to_byte_array(const uint8_t* bits, size_t n, uint8_t* out, size_t out_n)
{
    size_t i = 0, b = 0;
    uint8_t tmp = 0;
    for (size_t k = 0; k != n; ++k) {
        tmp |= (uint8_t) (bits[k] << (7 - b));
        if (++b == 8) { out[i] = tmp; tmp = 0; ++i; b = 0; }
    }
    if (i < out_n) out[i] = tmp;
}

/* ------------------------------------------------------------------------- */
/* frame decoder (M17FrameDecoder.h)                                         */
/* ------------------------------------------------------------------------- */

static void emit(m17p_demod* d, uint8_t type, int cost)
{
    d->dec_out_type = type;
    if (d->rec_n < d->rec_cap) {
        m17o_record* r = &d->rec[d->rec_n];
        memset(r, 0, sizeof(*r));
        r->stream = d->stream_id;
        r->seq = (uint32_t) d->rec_n;
        r->sample_pos = d->pos;
        r->viterbi_cost = cost;
        r->type = type;
        memcpy(r->payload, d->dec_payload, 26);
        memcpy(r->lsf, d->dec_lsf, 30);
    }
    d->rec_n += 1;
}

/* M17FrameDecoder.h:113-136 */
static void update_state(m17p_demod* d, const uint8_t* lsf_bits)
{
    if (lsf_bits[111]) {
        if (lsf_bits[109] != 0) d->dec_state = DEC_STREAM;
    } else {
        uint8_t packet_type = (uint8_t) ((lsf_bits[109] << 1) | lsf_bits[110]);
        d->dec_state = packet_type == 1 ? DEC_BASIC_PACKET : DEC_FULL_PACKET;
    }
}

/* M17FrameDecoder.h:154-178 */
static int decode_lsf(m17p_demod* d, const int8_t* buf)
{
    uint8_t bits[240];
    depuncture(buf, 368, d->dec_depunct, 488, g_p1, 61);
    d->viterbi_cost = viterbi_decode(d->dec_depunct, 488, bits, 240);
    to_byte_array(bits, 240, d->dec_lsf, 30);
    if (crc16(d->dec_lsf, 30) == 0) {
        update_state(d, bits);
        emit(d, M17O_LSF, (int) d->viterbi_cost);
        return RES_OK;
    }
    d->dec_lich_segments = 0;
    memset(d->dec_lsf, 0, 30);
    return RES_FAIL;
}

/* M17FrameDecoder.h:181-262 */
static int decode_lich(m17p_demod* d, const int8_t* buf)
{
    uint8_t* lich = d->dec_payload;
    memset(lich, 0, 6);
    size_t index = 0;
    for (size_t i = 0; i != 4; ++i) {
        uint32_t codeword = 0, decoded = 0;
        for (size_t j = 0; j != 24; ++j) {
            codeword <<= 1;
            codeword |= (buf[i * 24 + j] > 0);
        }
        if (!golay_decode(codeword, &decoded)) return RES_FAIL;
        decoded >>= 12;
        if (i & 1) {
            lich[index++] |= (uint8_t) (decoded >> 8);
            lich[index++] = (uint8_t) (decoded & 0xFF);
        } else {
            lich[index++] |= (uint8_t) (decoded >> 4);
            lich[index] = (uint8_t) ((decoded & 0x0F) << 4);
        }
    }
    emit(d, M17O_LICH, 0);

    uint8_t fragment_number = (uint8_t) ((lich[5] >> 5) & 7);
    if (fragment_number > 5) { d->viterbi_cost = (uint64_t) -1; return RES_INCOMPLETE; }
    memcpy(d->dec_lsf + fragment_number * 5, lich, 5);
    d->dec_lich_segments |= (uint8_t) (1 << fragment_number);
    if ((d->dec_lich_segments & 0x3F) != 0x3F) { d->viterbi_cost = (uint64_t) -1; return RES_INCOMPLETE; }
    if (crc16(d->dec_lsf, 30) == 0) {
        d->dec_lich_segments = 0;
        d->dec_state = DEC_STREAM;
        d->viterbi_cost = 0;
        emit(d, M17O_LSF, 0);
        return RES_OK;
    }
    d->viterbi_cost = 128;
    return RES_INCOMPLETE;
}

/* M17FrameDecoder.h:264-274.  402 does not divide into the P2 pattern: the
 * depuncture loop stops after consuming 368 inputs at output 400, so element
 * 401 keeps whatever an earlier LSF/packet decode left there. */
static int decode_bert(m17p_demod* d, const int8_t* buf)
{
    uint8_t bits[197];
    depuncture(buf, 368, d->dec_depunct, 402, g_p2, 12);
    d->viterbi_cost = viterbi_decode(d->dec_depunct, 402, bits, 197);
    to_byte_array(bits, 197, d->dec_payload, 25);
    emit(d, M17O_BERT, (int) d->viterbi_cost);
    return RES_OK;
}

/* M17FrameDecoder.h:276-289 */
static int decode_stream(m17p_demod* d, const int8_t* buf)
{
    uint8_t bits[144];
    depuncture(buf + 96, 272, d->dec_depunct, 296, g_p2, 12);
    d->viterbi_cost = viterbi_decode(d->dec_depunct, 296, bits, 144);
    to_byte_array(bits, 144, d->dec_payload, 18);
    emit(d, M17O_STREAM, (int) d->viterbi_cost);
    return RES_OK;
}

/* M17FrameDecoder.h:299-315 */
static int decode_packet(m17p_demod* d, const int8_t* buf, uint8_t type)
{
    uint8_t bits[206];
    depuncture(buf, 368, d->dec_depunct, 420, g_p3, 8);
    d->viterbi_cost = viterbi_decode(d->dec_depunct, 420, bits, 206);
    to_byte_array(bits, 206, d->dec_payload, 26);
    emit(d, type, (int) d->viterbi_cost);      /* callback result is true in the oracle */
    if (d->dec_payload[25] & 0x80) { d->dec_state = DEC_LSF; return RES_OK; }
    return RES_PACKET_INCOMPLETE;
}

/* M17FrameDecoder.h:353-392 */
static int frame_decode(m17p_demod* d, uint8_t sync_type, const int8_t* frame)
{
    int8_t tmp[368], buf[368];
    for (int i = 0; i != 368; ++i) tmp[i] = (int8_t) (frame[i] * g_dc[i]);   /* M17Randomizer.h:43-49 */
    for (int i = 0; i != 368; ++i) buf[i] = tmp[g_il[i]];                    /* PolynomialInterleaver.h:47-58 */
    switch (sync_type) {
    case SW_LSF:
        d->dec_state = DEC_LSF;
        return decode_lsf(d, buf);
    case SW_STREAM:
        switch (d->dec_state) {
        case DEC_LSF: return decode_lich(d, buf);
        case DEC_STREAM: return decode_stream(d, buf);
        default: d->dec_state = DEC_LSF;
        }
        break;
    case SW_PACKET:
        switch (d->dec_state) {
        case DEC_BASIC_PACKET: return decode_packet(d, buf, M17O_BASIC_PACKET);
        case DEC_FULL_PACKET: return decode_packet(d, buf, M17O_FULL_PACKET);
        default: d->dec_state = DEC_LSF;
        }
        break;
    case SW_BERT:
        d->dec_state = DEC_BERT;
        return decode_bert(d, buf);
    }
    return RES_FAIL;
}

/* ------------------------------------------------------------------------- */
/* float blocks                                                              */
/* ------------------------------------------------------------------------- */

/* FirFilter.h:28-43: newest sample times tap 0 first, strictly sequential. */
static float fir_step(m17p_demod* d, float input)
{
    d->fir_hist[d->fir_pos++] = input;
    if (d->fir_pos == 150) d->fir_pos = 0;
    float result = 0.0f;
    uint32_t index = d->fir_pos;
    for (int i = 0; i != 150; ++i) {
        index = index != 0 ? index - 1 : 149;
        result += d->fir_hist[index] * g_taps[i];
    }
    return result;
}

/* SlidingDFT.h:118-132 + DataCarrierDetect.h:53-58.  complex<float> multiply as
 * gcc emits it without -ffast-math: (ac - bd, ad + bc). */
static void dcd_step(m17p_demod* d, float sample)
{
    uint32_t index = d->dft_index;
    d->dft_index += 1;
    if (d->dft_index == 120) d->dft_index = 0;
    float delta = sample - d->dft_samples[index];
    for (int i = 0; i != 2; ++i) {
        float a = d->dft_re[i] + delta, b = d->dft_im[i];
        float c = g_sdft_c[2 * i], e = g_sdft_c[2 * i + 1];
        float re = a * c - b * e;
        float im = a * e + b * c;
        d->dft_re[i] = re;
        d->dft_im[i] = im;
    }
    d->dft_samples[index] = sample;
    d->dcd_level_1 += d->dft_re[0] * d->dft_re[0] + d->dft_im[0] * d->dft_im[0];   /* std::norm */
    d->dcd_level_2 += d->dft_re[1] * d->dft_re[1] + d->dft_im[1] * d->dft_im[1];
}

/* DataCarrierDetect.h:63-69 (ltrigger 0.1, htrigger 4.0 from M17Demodulator.h:149) */
static void dcd_update(m17p_demod* d)
{
    d->dcd_level = (float) ((double) d->dcd_level * 0.8 + 0.2 * (double) (d->dcd_level_1 / d->dcd_level_2));
    d->dcd_level_1 = 0.0f;
    d->dcd_level_2 = 0.0f;
    d->dcd_triggered = d->dcd_triggered ? d->dcd_level > 0.1f : d->dcd_level > 4.0f;
}

/* Correlator.h:43-49 + IirFilter.h:26-42 */
static void correlator_sample(m17p_demod* d, float value)
{
    float* h = d->iir_hist;
    h[2] = h[1]; h[1] = h[0];
    h[0] = fabsf(value);
    h[0] -= kIirA[1] * h[1];
    h[0] -= kIirA[2] * h[2];
    float result = 0;
    result += kIirB[0] * h[0];
    result += kIirB[1] * h[1];
    result += kIirB[2] * h[2];
    d->corr_limit = result;
    d->corr_buf[d->corr_pos] = value;
    d->corr_prev_pos = d->corr_pos;
    if (++d->corr_pos == 80) d->corr_pos = 0;
}

/* Correlator.h:51-64 */
static float correlate(const m17p_demod* d, const int8_t* sync)
{
    float result = 0.f;
    uint32_t pos = d->corr_prev_pos + 10;
    for (int i = 0; i != 8; ++i) {
        if (pos >= 80) pos -= 80;
        result += (float) sync[i] * d->corr_buf[pos];
        pos += 10;
    }
    return result;
}

static inline uint8_t corr_index(const m17p_demod* d) { return (uint8_t) (d->corr_prev_pos % 10); }

/* Correlator.h:81-114, including `avg = max + min / 2.` (:97) */
static void outer_symbol_levels(const m17p_demod* d, uint32_t sample_index, float* mn_out, float* mx_out)
{
    float min_sum = 0, max_sum = 0;
    size_t min_count = 0, max_count = 0;
    float min_level = d->corr_buf[sample_index], max_level = d->corr_buf[sample_index];
    for (uint32_t i = sample_index; i < 80; i += 10) {
        float v = d->corr_buf[i];
        min_level = v < min_level ? v : min_level;      /* std::min(min_level, v) */
        max_level = max_level < v ? v : max_level;      /* std::max(max_level, v) */
    }
    float avg = (float) ((double) max_level + (double) min_level / 2.);
    for (uint32_t i = sample_index; i < 80; i += 10) {
        float v = d->corr_buf[i];
        int high = v > avg, low = v < avg;
        max_sum += v * (float) high;
        min_sum += v * (float) low;
        max_count += (size_t) high;
        min_count += (size_t) low;
    }
    *mn_out = min_count > 0 ? min_sum / (float) min_count : min_level;
    *mx_out = max_count > 0 ? max_sum / (float) max_count : max_level;
}

/* Correlator.h:150-157 */
static float sync_triggered(const m17p_syncword* s, const m17p_demod* d)
{
    float limit_1 = d->corr_limit * s->mag1;
    float limit_2 = d->corr_limit * s->mag2;
    float value = correlate(d, s->word);
    return (value > limit_1 || value < limit_2) ? value : 0.0f;
}

/* Correlator.h:161-200 */
static uint8_t sync_step(m17p_syncword* s, const m17p_demod* d)
{
    float value = sync_triggered(s, d);
    if (value != 0) {
        if (!s->triggered) {
            for (int i = 0; i != 10; ++i) s->samples[i] = 0;
            s->triggered = 1;
        }
        s->samples[corr_index(d)] = value;
    } else if (s->triggered) {
        /* find_peak(value == 0) */
        s->triggered = 0;
        s->timing_index = 0;
        float peak_value = value;
        for (uint8_t i = 0; i != 10; ++i) {
            float f = s->samples[i];
            if (fabsf(f) > fabsf(peak_value)) { peak_value = f; s->timing_index = i; }
        }
        s->updated = peak_value > 0 ? 1 : -1;
    }
    return s->timing_index;
}

static int8_t sync_updated(m17p_syncword* s) { int8_t r = s->updated; s->updated = 0; return r; }

/* KalmanFilter.h:33-38 / :83-88 */
static void kf_reset(m17p_kalman* k, float z)
{
    k->x[0] = z; k->x[1] = 0.f;
    k->P[0][0] = 4.f; k->P[0][1] = 0.f; k->P[1][0] = 0.f; k->P[1][1] = 0.00000025f;
}

/*
 * KalmanFilter.h:40-65 (wrap = 1, modulo-10 innovation) and :90-107 (wrap = 0).
 * The reference evaluates this through blaze expression templates; the order
 * and precision below are read off the compiled reference (gcc 13 -O3, SSE2):
 *   x = F x ; P = (F P) F^T + Q                       all float, 2-term sums
 *   S = (H P) H^T + R  (float) ; sinv = 1.0 / (double) S
 *   y = z - H x (float)
 *   x_i += (float) ((double) (float)(P_i0 y + P_i1 (0 y)) * sinv)        float add
 *   P_ij = (float) ((double) P_ij - ((double)(float)(P_0j a_i) + (double)(float)((a_i 0) P_1j)) * sinv)
 *          with a_i = P_i0 * 1 + P_i1 * 0
 * H = [1 0], F = [[1 dt][0 1]]; the multiplications by 0 and 1 are kept so that
 * NaN/Inf propagate exactly as they do in the reference.
 */
static void kf_update(m17p_kalman* k, float z, uint64_t dt_count, int wrap)
{
    static const float Q[2][2] = {{6.25e-13f, 1.25e-12f}, {1.25e-12f, 2.50e-12f}};
    const float R = 0.5f, one = 1.f, zero = 0.f;
    float dt = (float) dt_count;
    float x0 = k->x[0] * one + k->x[1] * dt;
    float x1 = k->x[0] * zero + k->x[1] * one;
    float T00 = one * k->P[0][0] + dt * k->P[1][0];
    float T01 = one * k->P[0][1] + dt * k->P[1][1];
    float T10 = zero * k->P[0][0] + one * k->P[1][0];
    float T11 = zero * k->P[0][1] + one * k->P[1][1];
    float P00 = (one * T00 + dt * T01) + Q[0][0];
    float P01 = (zero * T00 + one * T01) + Q[0][1];
    float P10 = (one * T10 + dt * T11) + Q[1][0];
    float P11 = (zero * T10 + one * T11) + Q[1][1];
    float HP0 = one * P00 + zero * P10;
    float HP1 = one * P01 + zero * P11;
    float S = (HP0 * one + HP1 * zero) + R;
    double sinv = 1.0 / (double) S;

    if (wrap) {
        if ((double) (z - x0) < -5.0) z += 10.f;
        else if ((double) (z - x0) > 5.0) z -= 10.f;
    }
    float y = z - (x0 * one + x1 * zero);
    float hy0 = y * one, hy1 = y * zero;
    float k0 = hy0 * P00 + hy1 * P01;
    float k1 = hy0 * P10 + hy1 * P11;
    x0 = x0 + (float) ((double) k0 * sinv);
    x1 = x1 + (float) ((double) k1 * sinv);
    if (wrap) {
        while (x0 >= 10.f) x0 -= 10.f;
        while (x0 < 0) x0 += 10.f;
    }
    float a0 = P00 * one + P01 * zero;
    float a1 = P10 * one + P11 * zero;
    float a0h0 = a0 * one, a0h1 = a0 * zero, a1h0 = a1 * one, a1h1 = a1 * zero;
    float n00 = (float) ((double) P00 - ((double) (P00 * a0h0) + (double) (a0h1 * P10)) * sinv);
    float n01 = (float) ((double) P01 - ((double) (P01 * a0h0) + (double) (a0h1 * P11)) * sinv);
    float n10 = (float) ((double) P10 - ((double) (P00 * a1h0) + (double) (a1h1 * P10)) * sinv);
    float n11 = (float) ((double) P11 - ((double) (P01 * a1h0) + (double) (a1h1 * P11)) * sinv);
    k->x[0] = x0; k->x[1] = x1;
    k->P[0][0] = n00; k->P[0][1] = n01; k->P[1][0] = n10; k->P[1][1] = n11;
}

static int8_t wrap_index(int8_t s)
{
    s = s < 0 ? (int8_t) (s + 10) : s;
    s = s >= 10 ? (int8_t) (s - 10) : s;
    return s;
}

/* ClockRecovery.h:33-39 */
static void clock_reset(m17p_demod* d, float index)
{
    kf_reset(&d->clock_kf, index);
    d->clock_count = 0;
    d->clock_sample_index = (int8_t) index;
    d->clock_estimate = 0.f;
}

/* ClockRecovery.h:54-67 */
static void clock_update_index(m17p_demod* d, uint8_t index)
{
    kf_update(&d->clock_kf, (float) index, d->clock_count, 1);
    d->sample_estimate = d->clock_kf.x[0];
    d->clock_sample_index = wrap_index((int8_t) round((double) d->sample_estimate));
    d->clock_estimate = d->clock_kf.x[1];
    d->clock_count = 0;
}

/* ClockRecovery.h:76-88 */
static void clock_update(m17p_demod* d)
{
    double csw = fmod((double) (d->sample_estimate + d->clock_estimate * (float) d->clock_count), 10.0);
    if (csw < 0.) csw += 10;
    else if (csw >= 10) csw -= 10;
    d->clock_sample_index = wrap_index((int8_t) round(csw));
}

/* FreqDevEstimator.h:31-48 */
static void dev_update(m17p_demod* d, float minValue, float maxValue)
{
    kf_update(&d->dev_min, minValue, 192, 0);
    kf_update(&d->dev_max, maxValue, 192, 0);
    const float* mnf = d->dev_min.x;
    const float* mxf = d->dev_max.x;
    d->dev_offset = (float) ((double) (mxf[0] + mnf[0]) / 2.);
    d->dev_idev = (float) (6.0 / (double) (mxf[0] - mnf[0]));
    if (isnan(mnf[0]) || isnan(mnf[1]) || isnan(mxf[0]) || isnan(mxf[1])) d->dev_reset = 1;
    if (d->dev_reset) {
        d->dev_reset = 0;
        kf_reset(&d->dev_min, minValue);
        kf_reset(&d->dev_max, maxValue);
        d->dev_offset = (minValue + maxValue) / 2;
        d->dev_idev = (float) (6.0 / (double) (maxValue - minValue));
    }
}

/* SymbolEvm.h:31-51 + StandardDeviation.h:57-83 (alpha = float(1.0/184)) */
static void evm_update(m17p_demod* d, float sample)
{
    float e;
    if (sample > 2) e = sample - 3;
    else if (sample > 0) e = sample - 1;
    else if (sample > -2) e = sample + 1;
    else e = sample + 3;
    const float alpha = (float) (1.0 / 184);
    d->evm_S -= d->evm_S * alpha;
    d->evm_S += (e * e) * alpha;
}

/* Util.h:128-145 */
static void llr4(float sample, int8_t* a, int8_t* b)
{
    float t = (-3.0f < sample) ? sample : -3.0f;     /* std::max(MIN_VALUE, sample) */
    float s = (t < 3.0f) ? t : 3.0f;                 /* std::min(MAX_VALUE, t) */
    int lo = 0, len = 43;                            /* std::lower_bound: first thr >= s */
    while (len > 0) {
        int half = len >> 1;
        if (g_llr_thr[lo + half] < s) { lo = lo + half + 1; len = len - half - 1; }
        else len = half;
    }
    if (lo == 43) lo = 42;
    *a = g_llr_val[lo][0];
    *b = g_llr_val[lo][1];
}

/* ------------------------------------------------------------------------- */
/* demodulator (M17Demodulator.h)                                            */
/* ------------------------------------------------------------------------- */

static void set_sync(m17p_syncword* s, const int8_t w[8], float m1, float m2)
{
    memset(s, 0, sizeof(*s));
    memcpy(s->word, w, 8);
    s->mag1 = m1; s->mag2 = m2;
}

void m17p_init(m17p_demod* d)
{
    m17p_setup();
    memset(d, 0, sizeof(*d));
    /* M17Demodulator.h:154-157; one-sided words use numeric_limits<float>::lowest() */
    static const int8_t w_pre[8] = {+3, -3, +3, -3, +3, -3, +3, -3};
    static const int8_t w_lsf[8] = {+3, +3, +3, +3, -3, -3, +3, -3};
    static const int8_t w_pkt[8] = {3, -3, 3, 3, -3, -3, -3, -3};
    static const int8_t w_eot[8] = {+3, +3, +3, +3, +3, +3, -3, +3};
    set_sync(&d->preamble_sync, w_pre, 29.f, -3.402823466e+38f);
    set_sync(&d->lsf_sync, w_lsf, 31.f, -31.f);
    set_sync(&d->packet_sync, w_pkt, 31.f, -31.f);
    set_sync(&d->eot_sync, w_eot, 31.f, -3.402823466e+38f);
    kf_reset(&d->clock_kf, 0.f);
    kf_reset(&d->dev_min, 0.f);
    kf_reset(&d->dev_max, 0.f);
    d->dev_reset = 1;
    d->evm_S = 1.0f;                 /* StandardDeviation.h:60 */
    d->polarity = 1;
    d->demod_state = ST_UNLOCKED;
    d->sync_word_type = SW_LSF;
    d->dec_state = DEC_LSF;
    d->initializing = 1920;          /* M17Demodulator.h:659, per stream here */
}

/* M17Demodulator.h:232-241 */
static void update_values(m17p_demod* d, uint8_t index)
{
    float mn, mx;
    outer_symbol_levels(d, d->sample_index, &mn, &mx);
    dev_update(d, mn, mx);
    d->sync_sample_index = index;
}

/* M17Demodulator.h:274-286 with dcd_on (:243-257) / dcd_off (:259-265) */
static void update_dcd(m17p_demod* d)
{
    if (!d->dcd && d->dcd_triggered) {
        d->dcd = 1;
        if (d->demod_state == ST_UNLOCKED) {
            d->sync_count = 0;
            d->missing_sync_count = 0;
            memset(d->framer_buf, 0, sizeof(d->framer_buf));   /* framer.reset() */
            d->framer_index = 0;
            d->dec_state = DEC_LSF;                             /* decoder.reset() (frame_number unused) */
            d->evm_S = 0.0f;                                    /* evm.reset() */
        }
        d->need_clock_reset = 1;
    } else if (d->dcd && !d->dcd_triggered) {
        d->demod_state = ST_UNLOCKED;
        d->dcd = 0;
    }
}

/* M17Demodulator.h:288-342 */
static void do_unlocked(m17p_demod* d)
{
    if (d->missing_sync_count < 1920) {
        d->missing_sync_count += 1;
        uint8_t sync_index = sync_step(&d->preamble_sync, d);
        int8_t upd = sync_updated(&d->preamble_sync);
        if (upd) {
            d->sync_count = 0;
            d->missing_sync_count = 0;
            d->need_clock_reset = 1;
            d->dev_reset = 1;
            d->sample_index = sync_index;
            update_values(d, sync_index);
            d->demod_state = ST_LSF_SYNC;
        }
        return;
    }
    uint8_t sync_index = sync_step(&d->lsf_sync, d);
    int8_t upd = sync_updated(&d->lsf_sync);
    if (upd) {
        d->sync_count = 86;
        d->missing_sync_count = 0;
        d->need_clock_reset = 1;
        d->dev_reset = 1;
        d->sample_index = sync_index;
        update_values(d, sync_index);
        d->demod_state = ST_FRAME;
        d->sync_word_type = upd < 0 ? SW_STREAM : SW_LSF;
    }
    sync_index = sync_step(&d->packet_sync, d);
    upd = sync_updated(&d->packet_sync);
    if (upd < 0) {
        d->sync_count = 86;
        d->missing_sync_count = 0;
        d->need_clock_reset = 1;
        d->dev_reset = 1;
        d->sample_index = sync_index;
        update_values(d, sync_index);
        d->demod_state = ST_FRAME;
        d->sync_word_type = SW_BERT;
    }
}

/* M17Demodulator.h:349-411 */
static void do_lsf_sync(m17p_demod* d)
{
    if (corr_index(d) != d->sample_index) return;
    float sync_trig = sync_triggered(&d->preamble_sync, d);
    if ((double) sync_trig > 0.1) {
        d->need_clock_update = 1;
        d->sync_count += 1;
        return;
    }
    sync_trig = sync_triggered(&d->lsf_sync, d);
    float bert_trig = sync_triggered(&d->packet_sync, d);
    if (bert_trig < 0) {
        d->missing_sync_count = 0;
        d->sync_count = 86;
        d->need_clock_update = 1;
        update_values(d, d->sample_index);
        d->demod_state = ST_FRAME;
        d->sync_word_type = SW_BERT;
    } else if ((double) fabsf(sync_trig) > 0.1) {
        d->missing_sync_count = 0;
        d->sync_count = 86;
        d->need_clock_update = 1;
        update_values(d, d->sample_index);
        d->demod_state = ST_FRAME;
        d->sync_word_type = sync_trig > 0 ? SW_LSF : SW_STREAM;
    } else if (++d->missing_sync_count > 192) {
        if (d->sync_count >= 10) {
            d->missing_sync_count = 0;
            d->need_clock_update = 1;
        } else {
            d->sync_count = 0;
            d->demod_state = ST_UNLOCKED;
            d->missing_sync_count = 0;
            d->dcd_triggered = 0;            /* dcd.unlock() */
        }
    } else {
        update_values(d, d->sample_index);
    }
}

/* M17Demodulator.h:419-482 */
static void do_stream_sync(m17p_demod* d)
{
    d->sync_count += 1;
    if (d->sync_count < 78) return;
    if (sync_triggered(&d->eot_sync, d) > 0.1f) {       /* EOT_TRIGGER_LEVEL is a float constant */
        d->sync_word_type = SW_STREAM;
        d->demod_state = ST_FRAME;
        d->eot_flag = 1;
        d->missing_sync_count = 0;
        return;
    }
    uint8_t sync_index = sync_step(&d->lsf_sync, d);
    int8_t upd = sync_updated(&d->lsf_sync);
    if (upd < 0) {
        d->missing_sync_count = 0;
        update_values(d, sync_index);
        d->sync_word_type = SW_STREAM;
        d->demod_state = ST_SYNC_WAIT;
        d->eot_flag = 0;
    } else if (d->sync_count > 86) {
        if (d->viterbi_cost < 80) {
            if (!d->missing_sync_count) d->missing_sync_count = 1;
            d->sync_word_type = SW_STREAM;
            d->demod_state = ST_FRAME;
        } else if (d->eot_flag) {
            d->demod_state = ST_UNLOCKED;
            d->dcd_triggered = 0;
        } else if (d->missing_sync_count < 10) {
            d->missing_sync_count += 1;
            d->sync_word_type = SW_STREAM;
            d->demod_state = ST_FRAME;
        } else {
            d->demod_state = ST_UNLOCKED;
            d->dcd_triggered = 0;
        }
        d->eot_flag = 0;
    }
}

/* M17Demodulator.h:488-530 (packet, cost limit 60, any polarity) and :535-574
 * (BERT, cost limit 80, negative polarity only) */
static void do_packet_or_bert_sync(m17p_demod* d, int bert)
{
    d->sync_count += 1;
    if (d->sync_count < 78) return;
    uint8_t sync_index = sync_step(&d->packet_sync, d);
    int8_t upd = sync_updated(&d->packet_sync);
    uint8_t swt = bert ? SW_BERT : SW_PACKET;
    if (bert ? (upd < 0) : (upd != 0)) {
        d->missing_sync_count = 0;
        update_values(d, sync_index);
        d->sync_word_type = swt;
        d->demod_state = ST_SYNC_WAIT;
    } else if (d->sync_count > 86) {
        if (d->viterbi_cost < (bert ? 80u : 60u)) {
            if (!d->missing_sync_count) d->missing_sync_count = 1;
            d->sync_word_type = swt;
            d->demod_state = ST_FRAME;
        } else if (d->missing_sync_count < 10) {
            d->missing_sync_count += 1;
            d->sync_word_type = swt;
            d->demod_state = ST_FRAME;
        } else {
            d->demod_state = ST_UNLOCKED;
            d->dcd_triggered = 0;
        }
    }
}

/* M17Demodulator.h:582-593 */
static void do_sync_wait(m17p_demod* d)
{
    if (d->sync_count < 86) { d->sync_count += 1; return; }
    d->need_clock_update = 1;
    d->demod_state = ST_FRAME;
}

/* M17Demodulator.h:595-654 */
static void do_frame(m17p_demod* d, float filtered_sample)
{
    int diff = (int) d->sample_index - (int) corr_index(d);
    if (abs(diff) == 5) {
        clock_update(d);
        d->sample_index = (uint8_t) d->clock_sample_index;
        return;
    }
    if (corr_index(d) != d->sample_index) return;

    float sample = filtered_sample - d->dev_offset;
    sample *= d->dev_idev;
    sample *= (float) d->polarity;
    evm_update(d, sample);

    int8_t a, b;
    llr4(sample, &a, &b);
    d->framer_buf[d->framer_index++] = a;              /* M17Framer.h:42-53 */
    d->framer_buf[d->framer_index++] = b;
    if (d->framer_index != 368) return;
    d->framer_index = 0;

    d->sync_count = 0;
    frame_decode(d, d->sync_word_type, d->framer_buf);
    switch (d->dec_state) {
    case DEC_STREAM: d->demod_state = ST_STREAM_SYNC; break;
    case DEC_LSF: d->demod_state = ST_STREAM_SYNC; break;
    case DEC_BERT: d->demod_state = ST_BERT_SYNC; break;
    default: d->demod_state = ST_PACKET_SYNC; break;
    }
}

/* M17Demodulator.h:656-753 */
void m17p_sample(m17p_demod* d, float input)
{
    d->count++;
    dcd_step(d, input);

    if (d->initializing) {
        --d->initializing;
        correlator_sample(d, fir_step(d, input));      /* initialize(), :267-272 */
        d->count = 0;
        return;
    }

    if (!d->dcd) {
        if (d->count % 384 == 0) {
            update_dcd(d);
            dcd_update(d);
            d->count = 0;
        }
        return;
    }

    float filtered = fir_step(d, input);
    correlator_sample(d, filtered);

    if (corr_index(d) == 0) {
        if (d->need_clock_reset) {
            clock_reset(d, (float) d->sync_sample_index);
            d->need_clock_reset = 0;
            d->sample_index = d->sync_sample_index;
        } else if (d->need_clock_update) {
            clock_update_index(d, d->sync_sample_index);
            d->need_clock_update = 0;
        }
    }
    d->clock_count++;                                  /* clock_recovery(filtered) */

    switch (d->demod_state) {
    case ST_UNLOCKED: do_unlocked(d); break;
    case ST_LSF_SYNC: do_lsf_sync(d); break;
    case ST_STREAM_SYNC: do_stream_sync(d); break;
    case ST_PACKET_SYNC: do_packet_or_bert_sync(d, 0); break;
    case ST_BERT_SYNC: do_packet_or_bert_sync(d, 1); break;
    case ST_SYNC_WAIT: do_sync_wait(d); break;
    case ST_FRAME: do_frame(d, filtered); break;
    }

    if (d->count % 960 == 0) {
        update_dcd(d);
        d->count = 0;
        dcd_update(d);
    }
}

void m17p_capture(const m17p_demod* d, m17o_trace* t)
{
    memset(t, 0, sizeof(*t));
    t->filtered = d->corr_buf[d->corr_prev_pos];
    t->limit = d->corr_limit;
    t->dcd_level = d->dcd_level;
    t->dev_offset = d->dev_offset;
    t->dev_idev = d->dev_idev;
    t->clock_est = d->clock_estimate;
    t->sample_est = d->sample_estimate;
    t->evm = sqrtf(d->evm_S);
    t->sync_count = d->sync_count;
    t->missing_sync_count = d->missing_sync_count;
    t->viterbi_cost = (int) d->viterbi_cost;
    t->count = (uint32_t) d->count;
    t->demod_state = d->demod_state;
    t->sample_index = d->sample_index;
    t->sync_sample_index = d->sync_sample_index;
    t->clock_sample_index = (uint8_t) d->clock_sample_index;
    t->dcd = d->dcd;
    t->dcd_triggered = d->dcd_triggered;
    t->sync_word_type = d->sync_word_type;
    t->decoder_state = d->dec_state;
    t->corr_index = corr_index(d);
    t->need_clock_reset = d->need_clock_reset;
    t->need_clock_update = d->need_clock_update;
    t->polarity = (uint8_t) d->polarity;
}

/* apps/m17-demod.cpp:484-490 */
size_t m17port_demod(const int16_t* samples, size_t n, unsigned flags, uint32_t stream_id,
                     m17o_record* rec, size_t rec_cap, m17o_trace* trace)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    d->rec = rec; d->rec_cap = rec_cap; d->stream_id = stream_id;
    size_t total = n + (((flags & M17O_EOF_REPEAT) && n) ? 1 : 0);
    for (size_t i = 0; i != total; ++i) {
        int16_t sample = samples[i < n ? i : n - 1];
        if (flags & M17O_INVERT) sample = (int16_t) (sample * -1);
        d->pos = i;
        m17p_sample(d, (float) ((double) sample / 41067.0));
        if (trace) m17p_capture(d, &trace[i]);
    }
    size_t r = d->rec_n;
    free(d);
    return r;
}

int m17port_demod_batch(const int16_t* samples, size_t n_streams, size_t n_samples, size_t stride,
                        unsigned flags, m17o_record* rec, size_t rec_cap, uint64_t* counts,
                        m17o_trace* trace, int jobs_unused)
{
    (void) jobs_unused;
    for (size_t s = 0; s != n_streams; ++s)
        counts[s] = m17port_demod(samples + s * stride, n_samples, flags, (uint32_t) s, rec + s * rec_cap,
                                  rec_cap, trace ? trace + s * n_samples : NULL);
    return 0;
}

/* ------------------------------------------------------------------------- */
/* component taps                                                            */
/* ------------------------------------------------------------------------- */

uint16_t m17port_crc16(const uint8_t* data, size_t n) { m17p_setup(); return crc16(data, n); }
uint32_t m17port_golay_encode24(uint16_t data) { return golay_encode24(data); }
int m17port_golay_decode(uint32_t input, uint32_t* output) { m17p_setup(); return golay_decode(input, output); }
void m17port_llr(float sample, int8_t* out2) { m17p_setup(); llr4(sample, &out2[0], &out2[1]); }

void m17port_llr_table(float* thr, int8_t* pairs)
{
    m17p_setup();
    for (int i = 0; i != 43; ++i) { thr[i] = g_llr_thr[i]; pairs[2 * i] = g_llr_val[i][0]; pairs[2 * i + 1] = g_llr_val[i][1]; }
}

void m17port_fir(const float* in, float* out, size_t n)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    for (size_t i = 0; i != n; ++i) out[i] = fir_step(d, in[i]);
    free(d);
}

void m17port_rrc_taps(float* taps150) { m17p_setup(); memcpy(taps150, g_taps, sizeof(g_taps)); }
float m17port_scale(int16_t s) { return (float) ((double) s / 41067.0); }

size_t m17port_dcd(const float* in, size_t n, size_t period, float* level, uint8_t* trig)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    size_t k = 0;
    for (size_t i = 0; i != n; ++i) {
        dcd_step(d, in[i]);
        if ((i + 1) % period == 0) { dcd_update(d); level[k] = d->dcd_level; trig[k] = d->dcd_triggered; ++k; }
    }
    free(d);
    return k;
}

void m17port_sdft(const float* in, size_t n, float* out4)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    for (size_t i = 0; i != n; ++i) {
        dcd_step(d, in[i]);
        out4[4 * i] = d->dft_re[0]; out4[4 * i + 1] = d->dft_im[0];
        out4[4 * i + 2] = d->dft_re[1]; out4[4 * i + 3] = d->dft_im[1];
    }
    free(d);
}

void m17port_limit_iir(const float* in, float* out, size_t n)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    for (size_t i = 0; i != n; ++i) { correlator_sample(d, in[i]); out[i] = d->corr_limit; }
    free(d);
}

void m17port_correlator(const float* in, size_t n, uint8_t sample_index, float* corr4, float* mnmx2,
                        float* limit, uint8_t* index)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    for (size_t i = 0; i != n; ++i) correlator_sample(d, in[i]);
    corr4[0] = correlate(d, d->preamble_sync.word);
    corr4[1] = correlate(d, d->lsf_sync.word);
    corr4[2] = correlate(d, d->packet_sync.word);
    corr4[3] = correlate(d, d->eot_sync.word);
    outer_symbol_levels(d, sample_index, &mnmx2[0], &mnmx2[1]);
    *limit = d->corr_limit;
    *index = corr_index(d);
    free(d);
}

/* Correlator.h:43-64 per sample: out5[5 i + 0] = limit_ after sample i, [1..4] = correlate() with the preamble,
 * LSF / stream, packet / BERT and EOT sync words (M17Demodulator.h:154-157), correlator clocked on every sample. */
void m17port_correlate_trace(const float* in, size_t n, float* out5)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    for (size_t i = 0; i != n; ++i) {
        correlator_sample(d, in[i]);
        out5[5 * i + 0] = d->corr_limit;
        out5[5 * i + 1] = correlate(d, d->preamble_sync.word);
        out5[5 * i + 2] = correlate(d, d->lsf_sync.word);
        out5[5 * i + 3] = correlate(d, d->packet_sync.word);
        out5[5 * i + 4] = correlate(d, d->eot_sync.word);
    }
    free(d);
}

void m17port_clock(const uint8_t* ops, const uint8_t* args, const uint32_t* dt, size_t n,
                   uint8_t* sample_index, float* clock_est, float* sample_est)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    for (size_t i = 0; i != n; ++i) {
        d->clock_count += dt[i];
        switch (ops[i]) {
        case 0: clock_reset(d, (float) args[i]); break;
        case 1: clock_update_index(d, args[i]); break;
        default: clock_update(d); break;
        }
        sample_index[i] = (uint8_t) d->clock_sample_index;
        clock_est[i] = d->clock_estimate;
        sample_est[i] = d->sample_estimate;
    }
    free(d);
}

void m17port_freqdev(const float* mn, const float* mx, const uint8_t* reset_before, size_t n,
                     float* offset, float* idev)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    for (size_t i = 0; i != n; ++i) {
        if (reset_before[i]) d->dev_reset = 1;
        dev_update(d, mn[i], mx[i]);
        offset[i] = d->dev_offset;
        idev[i] = d->dev_idev;
    }
    free(d);
}

void m17port_evm(const float* in, size_t n, float* out)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    d->evm_S = 0.0f;
    for (size_t i = 0; i != n; ++i) { evm_update(d, in[i]); out[i] = sqrtf(d->evm_S); }
    free(d);
}

size_t m17port_frame_decoder(const int8_t* frames, const uint8_t* sync, size_t n_frames,
                             m17o_record* rec, size_t rec_cap, int64_t* costs, uint8_t* states,
                             uint8_t* results)
{
    m17p_demod* d = (m17p_demod*) malloc(sizeof(m17p_demod));
    m17p_init(d);
    d->rec = rec; d->rec_cap = rec_cap;
    for (size_t i = 0; i != n_frames; ++i) {
        d->pos = i;
        int r = frame_decode(d, sync[i], frames + i * 368);
        costs[i] = (int64_t) d->viterbi_cost;
        states[i] = d->dec_state;
        results[i] = (uint8_t) r;
    }
    size_t n = d->rec_n;
    free(d);
    return n;
}

size_t m17port_viterbi(int kind, const int8_t* in, uint8_t* out)
{
    static const size_t n_in[4] = {488, 296, 420, 402}, n_out[4] = {240, 144, 206, 197};
    m17p_setup();
    return (size_t) viterbi_decode(in, n_in[kind & 3], out, n_out[kind & 3]);
}

void m17port_tables(uint16_t* interleave368, int8_t* dc368, int8_t* p1_61, int8_t* p2_12, int8_t* p3_8)
{
    m17p_setup();
    memcpy(interleave368, g_il, sizeof(g_il));
    memcpy(dc368, g_dc, sizeof(g_dc));
    memcpy(p1_61, g_p1, 61); memcpy(p2_12, g_p2, 12); memcpy(p3_8, g_p3, 8);
}

void m17port_sdft_coeffs(float* re_im4) { m17p_setup(); memcpy(re_im4, g_sdft_c, sizeof(g_sdft_c)); }
