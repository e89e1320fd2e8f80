// m17b_tables.cpp -- host-side construction of the constant tables the kernels use.
//
// Everything here is computed the way the reference computes it at compile or
// start-up time, so that the uploaded constants are bit-identical:
//   RRC taps          M17Demodulator.h:79-118  (double literals narrowed to float)
//   SDFT coefficients SlidingDFT.h:85-95       (run-time std::exp in complex<float>)
//   LLR map           Util.h:63-104            (threshold by repeated float addition)
//   Viterbi input map M17Randomizer.h:29-49, PolynomialInterleaver.h:21-58,
//                     Trellis.h:17-40, Util.h:169-190, M17FrameDecoder.h:154-315
//   Golay syndromes   Golay24.h:102-177
//   CRC16 register    CRC16.h:21-34
#include "m17b_dev.cuh"

#include <cmath>
#include <complex>
#include <cstring>

namespace m17b {

namespace {

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

const uint8_t kDC[46] = {
    0xd6, 0xb5, 0xe2, 0x30, 0x82, 0xFF, 0x84, 0x62, 0xba, 0x4e, 0x96, 0x90, 0xd8, 0x98, 0xdd, 0x5d,
    0x0c, 0xc8, 0x52, 0x43, 0x91, 0x1d, 0xf8, 0x6e, 0x68, 0x2F, 0x35, 0xda, 0x14, 0xea, 0xcd, 0x76,
    0x19, 0x8d, 0xd5, 0x80, 0xd1, 0x33, 0x87, 0x13, 0x57, 0x18, 0x2d, 0x29, 0x78, 0xc3};

uint32_t golay_syndrome(uint32_t cw)
{
    cw &= 0xffffffu;
    for (int i = 0; i != 12; ++i) { if (cw & 1) cw ^= 0xC75; cw >>= 1; }
    return cw;     // 11-bit syndrome
}

// Depuncture pattern walk (Util.h:169-190) composed with the deinterleaver and
// the derandomiser sign: out[i] = 0 (punctured) or +-frame[k].
void make_vmap(int16_t* map, int n_out, int n_in, int in_offset, const int8_t* pat, int plen)
{
    int index = 0, pindex = 0, i = 0;
    for (; i != n_out && index < n_in; ++i) {
        if (!pat[pindex++]) {
            map[i] = 0;
        } else {
            unsigned j = unsigned(in_offset + index++);                 // index into the deinterleaved frame
            unsigned k = (45u * j + 92u * j * j) % 368u;                  // deinterleave: buf[j] = frame[k]
            int sign = ((kDC[k >> 3] >> (7 - (k & 7))) & 1) ? -1 : 1;    // derandomise: frame[k] *= dc[k]
            map[i] = int16_t(sign * int(k + 1));
        }
        if (pindex == plen) pindex = 0;
    }
    for (; i != n_out; ++i) map[i] = kMapStale;    // elements the reference's loop never writes
}

} // namespace

void build_tables(ConstTables& t, uint32_t* golay)
{
    std::memset(&t, 0, sizeof(t));
    for (int i = 0; i != kTaps; ++i)
        t.taps[i] = i == 149 ? 0.0f : float(i <= 74 ? kRrcHalf[i] : kRrcHalf[148 - i]);

    {   // SlidingDFT.h:85-95 (NSlidingDFT<float,48000,120,2>{2400, 3600})
        using C = std::complex<float>;
        const C j{0, 1};
        const float pi2 = float(M_PI * 2.0);
        const size_t freqs[2] = {2400, 3600};
        for (int i = 0; i != 2; ++i) {
            float k = float(freqs[i]) / float(48000);
            C c = std::exp(-j * pi2 * k);
            t.sdft[2 * i] = c.real();
            t.sdft[2 * i + 1] = c.imag();
        }
    }

    {   // Util.h:63-104
        const int8_t limit = 7;
        const float inc = float(1.0 / double(float(limit)));
        int8_t i = limit, j = limit;
        float k = float(-3.0 + double(inc));
        for (int index = 0; index != 43; ++index) {
            t.llr_thr[index] = k;
            t.llr_val[index][0] = i;
            t.llr_val[index][1] = j;
            if (double(k) + 1.0 < 0) { j--; if (j == 0) j = -1; if (j < -limit) j = -limit; }
            else if (double(k) - 1.0 < 0) { i--; if (i == 0) i = -1; if (i < -limit) i = -limit; }
            else { j++; if (j == 0) j = 1; if (j > limit) j = limit; }
            k = k + inc;
        }
    }

    int8_t p1[61], p2[12], p3[8];
    for (int i = 0, j = 2; i != 61; ++i) { if (i == j) { p1[i] = 0; j += 4; } else p1[i] = 1; }
    for (int i = 0; i != 12; ++i) p2[i] = i != 11;
    for (int i = 0; i != 8; ++i) p3[i] = i != 7;
    make_vmap(t.vmap[0], 488, 368, 0, p1, 61);     // LSF     M17FrameDecoder.h:156
    make_vmap(t.vmap[1], 296, 272, 96, p2, 12);    // stream  :278-281 (soft bits 96..367)
    make_vmap(t.vmap[2], 420, 368, 0, p3, 8);      // packet  :301
    make_vmap(t.vmap[3], 402, 368, 0, p2, 12);     // BERT    :266 (element 401 stays stale)

    for (unsigned q = 0; q != 96; ++q) {            // unpack_lich reads buffer[0..95], M17FrameDecoder.h:181-193
        unsigned k = (45u * q + 92u * q * q) % 368u;
        int sign = ((kDC[k >> 3] >> (7 - (k & 7))) & 1) ? -1 : 1;
        t.lichmap[q] = int16_t(sign * int(k + 1));
    }

    {   // CRC16.h:21-34
        uint16_t reg = 0xFFFF;
        for (int i = 0; i != 16; ++i) {
            uint16_t bit = reg & 1;
            if (bit) reg ^= 0x5935;
            reg >>= 1;
            if (bit) reg |= 0x8000;
        }
        t.crc_init = reg;
    }

    // Golay24.h:131-177: error pattern (<= 3 bits of 23) by 11-bit syndrome; the
    // code is perfect so all 2048 syndromes occur exactly once.
    for (int i = 0; i != 2048; ++i) golay[i] = 0;
    for (int a = 0; a != 23; ++a) {
        golay[golay_syndrome(1u << a)] = 1u << a;
        for (int b = a + 1; b != 23; ++b) {
            golay[golay_syndrome((1u << a) | (1u << b))] = (1u << a) | (1u << b);
            for (int c = b + 1; c != 23; ++c) {
                uint32_t v = (1u << a) | (1u << b) | (1u << c);
                golay[golay_syndrome(v)] = v;
            }
        }
    }
}

} // namespace m17b
