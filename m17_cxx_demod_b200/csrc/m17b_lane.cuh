// m17b_lane.cuh -- the per-stream ("lane") logic of the scan, search and decode kernels.
//
// One lane = one M17 stream = one CUDA thread.  Everything here is straight-line
// per-lane code with no warp intrinsics, so the very same source also compiles
// with g++ as a single-lane host build (tests/emu/, test infrastructure only)
// that replays the kernels' arithmetic on the CPU against the oracle.  Under nvcc
// the functions are __device__ only: the product has no CPU path.
//
//   scan_lane()    M17Demodulator::operator()                M17Demodulator.h:656-753
//   search_lane()  the sync-word search of do_unlocked()     M17Demodulator.h:288-342, Correlator.h:43-64,146-157
//   decode_lane()  M17FrameDecoder::operator()               M17FrameDecoder.h:353-392
//
// Structure of scan_lane (why it looks nothing like the reference's loop): the
// reference does everything once per sample, but only the correlator's limit
// biquad (Correlator.h:43-49) really is a per-sample recurrence of the clocked
// path (the carrier detect's sliding DFT has its own lock-step pre-pass,
// dcd_lane).  The state machine around it acts on a few samples per symbol at
// most, and the correlator's 80-sample buffer is nothing but the last 80 filter
// outputs, which the sample ring already holds.  A lane therefore walks aligned
// blocks of 8 samples (one 32-byte sector of the sample ring): the limit filter
// runs over the block in registers, and the reference's per-sample logic runs
// only on the samples the state machine can act on (quiet_samples() gives the
// distance in closed form); sync-word correlations are evaluated on demand, on
// the few samples of a frame's sync window, straight from the sample ring.
// The one state that evaluates correlations on EVERY sample -- ST_UNLOCKED with
// the carrier detect on -- is feed-forward once the limit filter has run, so
// it is not walked serially at all: the lane asks fixup_kernel for a search pass
// over the samples ahead (one thread per sample) and then jumps from trigger to
// trigger through the masks that pass leaves.
// The carrier-detect cadence (384 / 960 samples) and the 1920-sample start-up
// are multiples of 8 from sample 0, so mode changes (filter clocked or not)
// only ever happen at block ends.
#pragma once
#include <cstddef>

#include "m17b_dev.cuh"

#include <math.h>
#include <string.h>

#ifdef __CUDACC__
#define M17B_FN __device__ __forceinline__
#define M17B_FN_COLD __device__ __noinline__
#define M17B_TAB c_tab
#define M17B_GOLAY g_golay
#else
#define M17B_FN inline
#define M17B_FN_COLD inline
#define M17B_TAB ::m17b::h_tab
#define M17B_GOLAY ::m17b::h_golay
#endif

namespace m17b {

#ifdef __CUDACC__
M17B_FN float u2f(uint32_t v) { return __uint_as_float(v); }
M17B_FN uint32_t f2u(float v) { return __float_as_uint(v); }
M17B_FN float mul_rn(float a, float b) { return __fmul_rn(a, b); }
M17B_FN float add_rn(float a, float b) { return __fadd_rn(a, b); }
M17B_FN int popc(uint32_t v) { return __popc(v); }
M17B_FN void ctr_add(unsigned long long* p, unsigned long long v) { atomicAdd(p, v); }
M17B_FN unsigned int ctr_add(unsigned int* p, unsigned int v) { return atomicAdd(p, v); }
#else
extern ConstTables h_tab;
extern uint32_t h_golay[2048];
M17B_FN float u2f(uint32_t v) { float f; memcpy(&f, &v, 4); return f; }
M17B_FN uint32_t f2u(float f) { uint32_t v; memcpy(&v, &f, 4); return v; }
M17B_FN float mul_rn(float a, float b) { return a * b; }     // built with -ffp-contract=off
M17B_FN float add_rn(float a, float b) { return a + b; }
M17B_FN int popc(uint32_t v) { return __builtin_popcount(v); }
M17B_FN void ctr_add(unsigned long long* p, unsigned long long v) { *p += v; }
M17B_FN unsigned int ctr_add(unsigned int* p, unsigned int v) { unsigned int o = *p; *p += v; return o; }
#endif

#ifdef __CUDACC__
M17B_FN int first_set(uint32_t m) { return __ffs((int) m) - 1; }
M17B_FN void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }
M17B_FN void publish_fence() { __threadfence(); }
#else
M17B_FN int first_set(uint32_t m) { return __builtin_ffs((int) m) - 1; }
M17B_FN void prefetch_l2(const void*) {}
M17B_FN void publish_fence() {}
#endif

// ---------------------------------------------------------------------------
// Kalman filters
// ---------------------------------------------------------------------------

struct Kalman { float x0, x1, p00, p01, p10, p11; };

M17B_FN Kalman kf_load(const uint32_t* sb, int f)
{
    Kalman k;
    k.x0 = u2f(sb[(f + 0) * 32]); k.x1 = u2f(sb[(f + 1) * 32]);
    k.p00 = u2f(sb[(f + 2) * 32]); k.p01 = u2f(sb[(f + 3) * 32]);
    k.p10 = u2f(sb[(f + 4) * 32]); k.p11 = u2f(sb[(f + 5) * 32]);
    return k;
}

M17B_FN void kf_store(uint32_t* sb, int f, const Kalman& k)
{
    sb[(f + 0) * 32] = f2u(k.x0); sb[(f + 1) * 32] = f2u(k.x1);
    sb[(f + 2) * 32] = f2u(k.p00); sb[(f + 3) * 32] = f2u(k.p01);
    sb[(f + 4) * 32] = f2u(k.p10); sb[(f + 5) * 32] = f2u(k.p11);
}

// KalmanFilter.h:33-38 / :83-88
M17B_FN void kf_reset(Kalman& k, float z)
{
    k.x0 = z; k.x1 = 0.f; k.p00 = 4.f; k.p01 = 0.f; k.p10 = 0.f; k.p11 = 0.00000025f;
}

// KalmanFilter.h:40-65 (wrap) / :90-107.  The reference evaluates this through
// blaze expression templates; operation order and precision below are those of
// the compiled reference (float 2-term sums; the gain scaled in double and
// rounded to float before the state add; the covariance update in double).
// The multiplications by H = [1 0] and F = [[1 dt][0 1]] entries are kept so
// that NaN/Inf propagate identically.
M17B_FN void kf_update(Kalman& k, float z, float dt, bool wrap)
{
    const float Q00 = 6.25e-13f, Q01 = 1.25e-12f, Q10 = 1.25e-12f, Q11 = 2.50e-12f, R = 0.5f;
    const float one = 1.f, zero = 0.f;
    float x0 = k.x0 * one + k.x1 * dt;
    float x1 = k.x0 * zero + k.x1 * one;
    float T00 = one * k.p00 + dt * k.p10;
    float T01 = one * k.p01 + dt * k.p11;
    float T10 = zero * k.p00 + one * k.p10;
    float T11 = zero * k.p01 + one * k.p11;
    float P00 = (one * T00 + dt * T01) + Q00;
    float P01 = (zero * T00 + one * T01) + Q01;
    float P10 = (one * T10 + dt * T11) + Q10;
    float P11 = (zero * T10 + one * T11) + Q11;
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
    k.p00 = (float) ((double) P00 - ((double) (P00 * a0h0) + (double) (a0h1 * P10)) * sinv);
    k.p01 = (float) ((double) P01 - ((double) (P01 * a0h0) + (double) (a0h1 * P11)) * sinv);
    k.p10 = (float) ((double) P10 - ((double) (P00 * a1h0) + (double) (a1h1 * P10)) * sinv);
    k.p11 = (float) ((double) P11 - ((double) (P01 * a1h0) + (double) (a1h1 * P11)) * sinv);
    k.x0 = x0; k.x1 = x1;
}

M17B_FN int wrap_index(int v)      // ClockRecovery.h:61-62 on int8
{
    v = v < 0 ? v + 10 : v;
    v = v >= 10 ? v - 10 : v;
    return v;
}

// (int8_t) round(double(v)) for a float v in (-128, 128): round half away from zero.
// trunc and the subtraction are exact in float, so no double arithmetic is needed.
M17B_FN int round_half_away(float v)
{
    if (!(v == v)) return 0;            // NaN converts to 0 through the reference's int8_t cast
    float t = truncf(v);
    float fr = v - t;
    int r = (int) t;
    if (fr >= 0.5f) r += 1;
    else if (fr <= -0.5f) r -= 1;
    return r;
}

// ---------------------------------------------------------------------------
// Clocked-sample history.  The correlator's 80-sample buffer (Correlator.h:25-49)
// is not materialised: its content is the last 80 matched-filter outputs the
// demodulator CLOCKED, i.e. ring_y itself back to the position the filter last
// resumed at after a carrier-detect gap, and before that the snapshot taken when
// the gap began (corr_snap).
// ---------------------------------------------------------------------------

struct LlrTable;

struct LaneIo {
    const LlrTable* llr_tab;
    uint32_t* sb;          // &st[block][0][lane]
    float* sw_samples;     // + s, stride np
    int16_t* llr;          // + s, stride np
    float* snap;           // + s, stride np: matched-filter history frozen by a carrier-detect gap
    float* csnap;          // + s, stride np: correlator history frozen by a carrier-detect gap
    const float* ring_y;
    const int16_t* ring_x;
    const uint32_t* mask_ring;
    const float2* ckpt_ring;
    uint32_t np, lmask, s;
    int cls;               // walk class the lane was listed in (selects the work lists it appends to)
    bool split;            // M17B_STREAMING: the two classes have a chain of rounds each (m17b_submit), and this is the acquiring streams'
};

M17B_FN size_t ring_off(uint32_t p, uint32_t lmask, uint32_t np, uint32_t s)
{
    return ((size_t) ((p >> 5) & lmask) * np + s) * kLine + (p & (kLine - 1));
}

// index of the 8-sample block that contains p in mask_ring / ckpt_ring ([RL][Np][4])
M17B_FN size_t blk_off(uint32_t p, uint32_t lmask, uint32_t np, uint32_t s)
{
    return ((size_t) ((p >> 5) & lmask) * np + s) * 4 + ((p >> 3) & 3u);
}

// y of the sample clocked d steps before the one at ring position p; k = samples clocked since the
// correlator last resumed, counting the one at p (saturates at kTapsUsed: no gap within reach)
M17B_FN float y_back(const LaneIo& e, uint32_t p, int k, int d)
{
    if (d < k) return e.ring_y[ring_off(p - (uint32_t) d, e.lmask, e.np, e.s)];
    return e.csnap[(size_t) (79 - (d - k)) * e.np];
}

// Correlator.h:51-64: sum over i of sync[i] * buffer_[(prev_pos + 10 (i + 1)) % 80],
// i.e. the samples 70, 60, ..., 0 steps back, oldest first.
M17B_FN float correlate(const LaneIo& e, uint32_t p, int k, uint32_t word /* bit i: sync[i] = -3 */)
{
    float result = 0.f;
#pragma unroll
    for (int i = 0; i != 8; ++i) {
        float sy = ((word >> i) & 1) ? -3.f : 3.f;
        result = result + sy * y_back(e, p, k, 70 - 10 * i);
    }
    return result;
}

// sync words, bit i set <=> symbol i is -3 (M17Demodulator.h:154-157)
constexpr uint32_t kWordPreamble = 0xAA;   // +3,-3,+3,-3,+3,-3,+3,-3
constexpr uint32_t kWordLsf = 0xB0;        // +3,+3,+3,+3,-3,-3,+3,-3
constexpr uint32_t kWordPacket = 0xF2;     // +3,-3,+3,+3,-3,-3,-3,-3
constexpr uint32_t kWordEot = 0x40;        // +3,+3,+3,+3,+3,+3,-3,+3
constexpr float kLowest = -3.402823466e+38f;   // numeric_limits<float>::lowest(), Correlator.h:143

// Correlator::sample's limit filter: BaseIirFilter<float,3> (IirFilter.h:26-42) with the coefficients of
// Correlator.h:38-39 on |filtered sample|.  h1 / h2 = the filter's history, newest first.
struct Biquad { float h1, h2; };

M17B_FN float limit_step(Biquad& b, float y)
{
    float h0 = fabsf(y);
    h0 -= -1.98148851f * b.h1;
    h0 -= 0.98165828f * b.h2;
    float r = 0.f;
    r += 4.24433681e-05f * h0;
    r += 8.48867363e-05f * b.h1;
    r += 4.24433681e-05f * b.h2;
    b.h2 = b.h1; b.h1 = h0;
    return r;
}

// SyncWord::triggered (Correlator.h:150-157) reduced to its decision, for the sync words of
// M17Demodulator.h:154-157: bit w of the result <=> triggered() returns non-zero for word w
// (0 preamble, 1 LSF / stream, 2 packet / BERT, 3 EOT) given limit_ and the four correlations.
M17B_FN uint32_t trigger_bits(float limit, float vp, float vl, float vk, float ve)
{
    const float l29 = limit * 29.f, l31 = limit * 31.f, lm31 = limit * -31.f, llow = limit * kLowest;
    uint32_t b = 0;
    b |= ((vp > l29 || vp < llow) && vp != 0.f) ? 1u : 0u;
    b |= ((vl > l31 || vl < lm31) && vl != 0.f) ? 2u : 0u;
    b |= ((vk > l31 || vk < lm31) && vk != 0.f) ? 4u : 0u;
    b |= ((ve > l31 || ve < llow) && ve != 0.f) ? 8u : 0u;
    return b;
}

// the correlation of Correlator.h:51-64 from the eight samples 70, 60, ..., 0 steps back (oldest first)
M17B_FN float corr8(const float (&y8)[8], uint32_t word)
{
    float result = 0.f;
#pragma unroll
    for (int i = 0; i != 8; ++i) {
        float sy = ((word >> i) & 1) ? -3.f : 3.f;
        result = result + sy * y8[i];
    }
    return result;
}

// Correlator.h:150-157
M17B_FN float sync_triggered(const LaneIo& e, uint32_t p, int k, float limit, uint32_t word, float m1, float m2)
{
    float limit_1 = limit * m1;
    float limit_2 = limit * m2;
    float value = correlate(e, p, k, word);
    return (value > limit_1 || value < limit_2) ? value : 0.0f;
}

// Correlator.h:161-207: SyncWord::operator() followed by updated(), given triggered()'s value.  `f` is
// the packed {timing_index:4, triggered:1}; returns new f | (updated & 3) << 8
// (updated: 0, 1, or 3 for -1).  samples = SyncWord::samples_ column, stride np.
M17B_FN_COLD uint32_t sync_step(float* samples, size_t np, float value, int idx10, uint32_t f)
{
    uint32_t timing_index = f & 0xF;
    bool trig = (f >> 4) & 1;
    uint32_t updated = 0;
    if (value != 0) {
        if (!trig) {
            for (int i = 0; i != 10; ++i) samples[(size_t) i * np] = 0.f;
            trig = true;
        }
        samples[(size_t) idx10 * np] = value;
    } else if (trig) {
        trig = false;
        timing_index = 0;
        float peak = value;
        for (int i = 0; i != 10; ++i) {
            float v = samples[(size_t) i * np];
            if (fabsf(v) > fabsf(peak)) { peak = v; timing_index = i; }
        }
        updated = peak > 0 ? 1u : 3u;
    }
    return timing_index | ((uint32_t) trig << 4) | (updated << 8);
}

// M17Demodulator.h:232-241: Correlator::outer_symbol_levels (Correlator.h:81-114,
// incl. `avg = max + min / 2.`) + FreqDevEstimator::update (FreqDevEstimator.h:31-48).
// pp = logical buffer position (0..79) of the sample being handled, p its ring position, k the
// samples clocked since the last resume.  Returns (offset, idev); dev.reset_ is always false afterwards.
struct F2 { float x, y; };

// y8 (may be null): the samples 70, 60, ..., 0 steps back, when the caller holds them and pp = si (mod 10).
M17B_FN_COLD F2 update_values(uint32_t* sb, const LaneIo& e, int si, int pp, uint32_t p, int k, bool dev_reset, const float* y8)
{
    float min_sum = 0, max_sum = 0;
    int min_count = 0, max_count = 0;
    float lv[8];
    int n = 0;
    for (int i = si; i < 80; i += 10) {
        int d = pp - i;
        d = d < 0 ? d + 80 : d;
        lv[n++] = y8 ? y8[7 - d / 10] : y_back(e, p, k, d);
    }
    float min_level = lv[0], max_level = lv[0];
    for (int i = 0; i != n; ++i) {
        float v = lv[i];
        min_level = v < min_level ? v : min_level;
        max_level = max_level < v ? v : max_level;
    }
    float avg = (float) ((double) max_level + (double) min_level / 2.);
    for (int i = 0; i != n; ++i) {
        float v = lv[i];
        bool high = v > avg, low = v < avg;
        max_sum += v * (high ? 1.f : 0.f);
        min_sum += v * (low ? 1.f : 0.f);
        max_count += high;
        min_count += low;
    }
    float minValue = min_count > 0 ? min_sum / (float) min_count : min_level;
    float maxValue = max_count > 0 ? max_sum / (float) max_count : max_level;

    Kalman mn = kf_load(sb, F_DMIN_X0), mx = kf_load(sb, F_DMAX_X0);
    kf_update(mn, minValue, 192.f, false);
    kf_update(mx, maxValue, 192.f, false);
    float offset = (float) ((double) (mx.x0 + mn.x0) / 2.);
    float idev = (float) (6.0 / (double) (mx.x0 - mn.x0));
    if (isnan(mn.x0) || isnan(mn.x1) || isnan(mx.x0) || isnan(mx.x1)) dev_reset = true;
    if (dev_reset) {
        kf_reset(mn, minValue);
        kf_reset(mx, maxValue);
        offset = (minValue + maxValue) / 2.f;
        idev = (float) (6.0 / (double) (maxValue - minValue));
    }
    kf_store(sb, F_DMIN_X0, mn);
    kf_store(sb, F_DMAX_X0, mx);
    F2 r; r.x = offset; r.y = idev;
    return r;
}

// ClockRecovery.h:54-67: Kalman update with the sync word's index; returns
// (sample_estimate_, clock_estimate_).
M17B_FN_COLD F2 clock_update_index(uint32_t* sb, int ssi, uint32_t ck_count)
{
    Kalman k = kf_load(sb, F_CK_X0);
    kf_update(k, (float) ssi, (float) ck_count, true);
    kf_store(sb, F_CK_X0, k);
    F2 r; r.x = k.x0; r.y = k.x1;
    return r;
}

// apps/m17-demod.cpp:489: float(double(sample) / 41067.0)
#ifdef __CUDACC__
M17B_FN float scale_i16(int v)
{
    // q = s*r, one Newton correction with two FMAs gives the correctly rounded float quotient, which
    // equals the reference's double-rounded value for all 65536 inputs (tests/test_gpu_parity.py)
    const float r = 0x1.988864p-16f;       // RN(1 / 41067)
    float sf = (float) v;
    float q = __fmul_rn(sf, r);
    float e = __fmaf_rn(-q, 41067.0f, sf);
    return __fmaf_rn(e, r, q);
}
#else
M17B_FN float scale_i16(int v) { return (float) ((double) v / 41067.0); }
#endif

// FIR history snapshot when the carrier detect gates the filter off
// (M17Demodulator.h:675-689: demod_filter is not clocked while !dcd_).  snap[j] =
// the j-th most recent input the filter has seen; k = samples clocked since the
// last resume (>= 149 unless the carrier dropped again at once).
M17B_FN_COLD void take_snapshot(float* snap, const int16_t* ring_x, uint32_t np, uint32_t lmask,
                                uint32_t s, uint32_t pos, int k)
{
    if (k < kTapsUsed)
        for (int j = kTapsUsed - 1; j >= k; --j) snap[(size_t) j * np] = snap[(size_t) (j - k) * np];
    int n = k < kTapsUsed ? k : kTapsUsed;
    for (int j = 0; j != n; ++j) snap[(size_t) j * np] = scale_i16(ring_x[ring_off(pos - (uint32_t) j, lmask, np, s)]);
}

// Util.h:128-145 (llr<float,4>) -> packed (second << 8 | first) int8 pair.  The reference
// binary-searches 43 thresholds spaced ~1/7 apart; the same lower_bound index is found here
// from an arithmetic first guess corrected against the thresholds themselves (a copy of the
// table in shared memory: lanes look up different entries).
struct LlrTable { float thr[48]; uint32_t val[44]; };      // thr[i + 2] = threshold i; -inf below, +inf above

// entry `i` (0..47) of the padded threshold column
M17B_FN float llr_thr_padded(const float* thr43, int i)
{
    return i < 2 ? -__builtin_huge_valf() : i < 45 ? thr43[i - 2] : __builtin_huge_valf();
}

M17B_FN uint32_t llr4(const LlrTable& tab, float sample)
{
    float t = (-3.0f < sample) ? sample : -3.0f;      // std::max(MIN_VALUE, sample)
    float v = (t < 3.0f) ? t : 3.0f;                  // std::min(MAX_VALUE, t)
    int g = (int) ((v + 3.0f) * 7.0f);                // about the number of thresholds below v
    g = g < 0 ? 0 : (g > 43 ? 43 : g);
    // lower_bound = the number of thresholds below v (they increase): four independent look-ups around
    // the guess settle it when it lies in g-1 .. g+1, which it does but for rounding at the far ends
    const float* w = tab.thr + g;                     // thresholds g-2 .. g+1
    const float a = w[0], b = w[1], c = w[2], d = w[3];
    const int n = (int) (a < v) + (int) (b < v) + (int) (c < v) + (int) (d < v);
    if (n >= 1 && n <= 3) g = g - 2 + n;
    else {
        while (g > 0 && !(tab.thr[g + 1] < v)) --g;   // first index with !(thr < v)
        while (g < 43 && tab.thr[g + 2] < v) ++g;
    }
    if (g == 43) g = 42;
    return tab.val[g];
}

// ---------------------------------------------------------------------------
// scan
// ---------------------------------------------------------------------------

// One 8-sample block of ring_y as a lane holds it while it walks the block.
struct Blk {
    float y[8];
    Biquad start;          // limit-filter history before block sample jb
    uint32_t base;         // ring position of y[0]
    int jb;                // first sample of the block this pass clocks
};

M17B_FN float y_sel(const float (&y)[8], int j)
{
    const float a = (j & 1) ? y[1] : y[0], b = (j & 1) ? y[3] : y[2], c = (j & 1) ? y[5] : y[4], d = (j & 1) ? y[7] : y[6];
    const float ab = (j & 2) ? b : a, cd = (j & 2) ? d : c;
    return (j & 4) ? cd : ab;
}

// limit_ after block sample j (Correlator.h:45): the block's filter steps jb .. j again
M17B_FN float limit_at(const Blk& b, int j)
{
    Biquad q = b.start;
    float r = 0.f;
#pragma unroll
    for (int i = 0; i != 8; ++i)
        if (i >= b.jb && i <= j) r = limit_step(q, b.y[i]);
    return r;
}

// Per-lane working set of the scan.
struct Scan {
    float level;                               // DCD level (the sliding DFT and its energy sums live in dcd_lane)
    Biquad bq;                                 // limit filter history (Correlator::sample_filter)
    float idev, offset, evm;
    float ck_est, ck_sest;
    uint32_t ck_count;
    int count, init;
    int corr_pos, i10, framer_index, gap_k;    // logical slot (0..79) the next clocked sample goes to, and that modulo 10
    int pp;                                    // logical slot of the sample being handled
    uint32_t p;                                // its ring position
    int k;                                     // samples clocked since the last resume, counting it (<= kTapsUsed)
    int state, swt, si, ssi, csi;
    bool dcd, dcd_trig, need_reset, need_update, dev_reset, eot, pending;
    int sync_count, missing;
    uint32_t vcost;
    uint32_t sw;                               // packed SyncWord state, 8 bits per word
    // where the sample being handled gets its sync-word trigger decisions from
    float limit;                               // block walk: limit_ after it, recomputed from the block's samples when first asked for
    bool have_limit;
    float y8[8];                               // block walk: the samples its correlations are made of (70, 60, ..., 0 steps back), loaded when first asked for
    bool have_y8;
    uint32_t mbits;                            // search walk: the search pass's trigger bits of this sample (bit w = sync word w)
};

// SyncWord::triggered() for the sample being handled, block walk
M17B_FN float sync_value(Scan& z, const LaneIo& e, const Blk& b, uint32_t word, float m1, float m2)
{
    if (!z.have_limit) {
        z.limit = limit_at(b, (int) (z.p - b.base));
        z.have_limit = true;
    }
    if (!z.have_y8) {                                  // every sync word of this sample correlates with the same eight
        if (z.k > 70) {
#pragma unroll
            for (int i = 0; i != 8; ++i) z.y8[i] = e.ring_y[ring_off(z.p - (uint32_t) (70 - 10 * i), e.lmask, e.np, e.s)];
        } else {
#pragma unroll
            for (int i = 0; i != 8; ++i) z.y8[i] = y_back(e, z.p, z.k, 70 - 10 * i);
        }
        z.have_y8 = true;
    }
    // Correlator.h:150-157
    const float limit_1 = z.limit * m1, limit_2 = z.limit * m2;
    const float value = corr8(z.y8, word);
    return (value > limit_1 || value < limit_2) ? value : 0.0f;
}

// ... search walk (w: 0 preamble, 1 LSF / stream, 2 packet / BERT): the decision is the search pass's, the
// value the same correlation
M17B_FN float sync_value_masked(const Scan& z, const LaneIo& e, int w, uint32_t word)
{
    return ((z.mbits >> w) & 1u) ? correlate(e, z.p, z.k, word) : 0.f;
}

#define M17B_UPDATE_VALUES(index_)                                                  \
    do {                                                                            \
        /* (a copy: a pointer into z would put the whole working set into local memory) */ \
        float y8c_[8];                                                              \
        const bool use8_ = z.have_y8 && (z.pp - z.si + 80) % 10 == 0;               \
        if (use8_) { _Pragma("unroll") for (int i_ = 0; i_ != 8; ++i_) y8c_[i_] = z.y8[i_]; } \
        F2 od_ = update_values(e.sb, e, z.si, z.pp, z.p, z.k, z.dev_reset, use8_ ? y8c_ : nullptr); \
        z.offset = od_.x; z.idev = od_.y; z.dev_reset = false;                      \
        z.ssi = (index_);                                                           \
    } while (0)

M17B_FN int do_sync_step(Scan& z, const LaneIo& e, int k, float value /* SyncWord::triggered() */, int* timing)
{
    uint32_t f = (z.sw >> (8 * k)) & 0xFF;
    uint32_t r = sync_step(e.sw_samples + (size_t) (10 * k) * e.np, e.np, value, z.pp % 10, f);
    z.sw = (z.sw & ~(0xFFu << (8 * k))) | ((r & 0xFF) << (8 * k));
    *timing = (int) (r & 0xF);
    uint32_t u = (r >> 8) & 3;
    return u == 0 ? 0 : (u == 1 ? 1 : -1);
}

// DataCarrierDetect.h:63-69 (ltrigger 0.1, htrigger 4.0)
// l1 / l2 = the energy sums of the two sliding-DFT bins over the window that ends here (dcd_sums)
M17B_FN void dcd_update(Scan& z, float l1, float l2)
{
    z.level = (float) ((double) z.level * 0.8 + 0.2 * (double) (l1 / l2));
    z.dcd_trig = z.dcd_trig ? z.level > 0.1f : z.level > 4.0f;
}

// M17Demodulator.h:274-286 (update_dcd, dcd_on :243-257, dcd_off :259-265)
M17B_FN void update_dcd(Scan& z, const LaneIo& e)
{
    if (!z.dcd && z.dcd_trig) {
        z.dcd = true;
        if (z.state == ST_UNLOCKED) {
            z.sync_count = 0;
            z.missing = 0;
            z.framer_index = 0;                                         // framer.reset()
            e.sb[F_DEC * 32] = (e.sb[F_DEC * 32] & ~0xFFu) | DEC_LSF;    // decoder.reset()
            z.evm = 0.f;                                                // evm.reset()
        }
        z.need_reset = true;
        z.gap_k = 0;                                                    // filter resumes on stale history
    } else if (z.dcd && !z.dcd_trig) {
        z.state = ST_UNLOCKED;
        z.dcd = false;
    }
}

// M17Demodulator.h:288-342 (search walk only: the trigger decisions are the search pass's)
M17B_FN void do_unlocked(Scan& z, const LaneIo& e)
{
    int timing;
    if (z.missing < 1920) {
        z.missing += 1;
        int upd = do_sync_step(z, e, 0, sync_value_masked(z, e, 0, kWordPreamble), &timing);
        if (upd) {
            z.sync_count = 0;
            z.missing = 0;
            z.need_reset = true;
            z.dev_reset = true;
            z.si = timing;
            M17B_UPDATE_VALUES(timing);
            z.state = ST_LSF_SYNC;
        }
        return;
    }
    int upd = do_sync_step(z, e, 1, sync_value_masked(z, e, 1, kWordLsf), &timing);
    if (upd) {
        z.sync_count = 86;
        z.missing = 0;
        z.need_reset = true;
        z.dev_reset = true;
        z.si = timing;
        M17B_UPDATE_VALUES(timing);
        z.state = ST_FRAME;
        z.swt = upd < 0 ? SW_STREAM : SW_LSF;
    }
    upd = do_sync_step(z, e, 2, sync_value_masked(z, e, 2, kWordPacket), &timing);
    if (upd < 0) {
        z.sync_count = 86;
        z.missing = 0;
        z.need_reset = true;
        z.dev_reset = true;
        z.si = timing;
        M17B_UPDATE_VALUES(timing);
        z.state = ST_FRAME;
        z.swt = SW_BERT;
    }
}

// M17Demodulator.h:349-411 (only called on samples with index() == sample_index)
M17B_FN void do_lsf_sync(Scan& z, const LaneIo& e, const Blk& b)
{
    float trig = sync_value(z, e, b, kWordPreamble, 29.f, kLowest);
    if ((double) trig > 0.1) {
        z.need_update = true;
        z.sync_count += 1;
        return;
    }
    trig = sync_value(z, e, b, kWordLsf, 31.f, -31.f);
    float bert = sync_value(z, e, b, kWordPacket, 31.f, -31.f);
    if (bert < 0) {
        z.missing = 0;
        z.sync_count = 86;
        z.need_update = true;
        M17B_UPDATE_VALUES(z.si);
        z.state = ST_FRAME;
        z.swt = SW_BERT;
    } else if ((double) fabsf(trig) > 0.1) {
        z.missing = 0;
        z.sync_count = 86;
        z.need_update = true;
        M17B_UPDATE_VALUES(z.si);
        z.state = ST_FRAME;
        z.swt = trig > 0 ? SW_LSF : SW_STREAM;
    } else if (++z.missing > 192) {
        if (z.sync_count >= 10) {
            z.missing = 0;
            z.need_update = true;
        } else {
            z.sync_count = 0;
            z.state = ST_UNLOCKED;
            z.missing = 0;
            z.dcd_trig = false;              // dcd.unlock()
        }
    } else {
        M17B_UPDATE_VALUES(z.si);
    }
}

// M17Demodulator.h:419-482 past the `sync_count < MIN_SYNC_COUNT` gate
M17B_FN void do_stream_sync(Scan& z, const LaneIo& e, const Blk& b)
{
    if (sync_value(z, e, b, kWordEot, 31.f, kLowest) > 0.1f) {
        z.swt = SW_STREAM;
        z.state = ST_FRAME;
        z.eot = true;
        z.missing = 0;
        return;
    }
    int timing;
    int upd = do_sync_step(z, e, 1, sync_value(z, e, b, kWordLsf, 31.f, -31.f), &timing);
    if (upd < 0) {
        z.missing = 0;
        M17B_UPDATE_VALUES(timing);
        z.swt = SW_STREAM;
        z.state = ST_SYNC_WAIT;
        z.eot = false;
    } else if (z.sync_count > 86) {
        if (z.vcost < 80u) {
            if (!z.missing) z.missing = 1;
            z.swt = SW_STREAM;
            z.state = ST_FRAME;
        } else if (z.eot) {
            z.state = ST_UNLOCKED;
            z.dcd_trig = false;
        } else if (z.missing < 10) {
            z.missing += 1;
            z.swt = SW_STREAM;
            z.state = ST_FRAME;
        } else {
            z.state = ST_UNLOCKED;
            z.dcd_trig = false;
        }
        z.eot = false;
    }
}

// M17Demodulator.h:488-530 (packet) and :535-574 (BERT) past the MIN_SYNC_COUNT gate
M17B_FN void do_packet_or_bert_sync(Scan& z, const LaneIo& e, const Blk& b, bool bert)
{
    int timing;
    int upd = do_sync_step(z, e, 2, sync_value(z, e, b, kWordPacket, 31.f, -31.f), &timing);
    int swt = bert ? SW_BERT : SW_PACKET;
    if (bert ? (upd < 0) : (upd != 0)) {
        z.missing = 0;
        M17B_UPDATE_VALUES(timing);
        z.swt = swt;
        z.state = ST_SYNC_WAIT;
    } else if (z.sync_count > 86) {
        if (z.vcost < (bert ? 80u : 60u)) {
            if (!z.missing) z.missing = 1;
            z.swt = swt;
            z.state = ST_FRAME;
        } else if (z.missing < 10) {
            z.missing += 1;
            z.swt = swt;
            z.state = ST_FRAME;
        } else {
            z.state = ST_UNLOCKED;
            z.dcd_trig = false;
        }
    }
}

// ClockRecovery.h:76-88: the sample index dead-reckoned from the clock estimate, `count` samples after the
// last update (clock_recovery.update() without an index, called half a symbol away from the symbol instant)
M17B_FN int clock_dead_reckon(float sest, float est, uint32_t count)
{
    // fmod(v, 10) is v itself for |v| < 10 and v -+ 10 (exact) for |v| < 20; libm otherwise
    const float v = sest + est * (float) count;
    const float av = fabsf(v);
    float csw;
    if (av < 10.f) csw = v;
    else if (av < 20.f) csw = v - (v < 0 ? -10.0f : 10.0f);
    else if (av < 16777216.f) {
        // fmod exactly, without libm's loop: v is a multiple of its ulp (>= 2^-19 here), so v / 10 is either an
        // integer or further from one than the division's rounding error, and v - 10 trunc(v / 10) is exact in double
        const double dv = (double) v;
        csw = (float) (dv - 10.0 * trunc(dv / 10.0));
    }
    else csw = (float) fmod((double) v, 10.0);      // exact: |result| < 10 has no more bits than v
    // `if (csw < 0.) csw += 10; else if (csw >= 10) csw -= 10;` in double, then round:
    // csw + 10 can need more than 24 bits, so keep that one step in double
    int r;
    if (csw < 0.f) r = (int) (int8_t) round((double) csw + 10.0);
    else r = round_half_away(csw);
    return wrap_index(r);
}

// One symbol of a frame (M17Demodulator.h:606-654 up to the decoder call): normalise, SymbolEvm, 4-FSK soft
// decision, framer.  Sets z.pending when the symbol completes the frame.
M17B_FN void frame_symbol(Scan& z, const LaneIo& e, float y)
{
    float sample = y - z.offset;
    sample *= z.idev;                                   // polarity is +1 and never changes
    // SymbolEvm.h:31-51 + StandardDeviation.h:68-72
    float err;
    if (sample > 2) err = sample - 3;
    else if (sample > 0) err = sample - 1;
    else if (sample > -2) err = sample + 1;
    else err = sample + 3;
    const float alpha = (float) (1.0 / 184);
    z.evm -= z.evm * alpha;
    z.evm += (err * err) * alpha;

    uint32_t pair = llr4(*e.llr_tab, sample);
    e.llr[(size_t) (z.framer_index >> 1) * e.np] = (int16_t) pair;      // M17Framer.h:42-53
    z.framer_index += 2;
    if (z.framer_index != 368) return;
    z.framer_index = 0;
    z.sync_count = 0;
    z.pending = true;                                   // decoder(...) runs in decode_kernel
    z.state = ST_TBD;
}

// M17Demodulator.h:595-654 up to the decoder call (which decode_kernel performs)
M17B_FN void do_frame(Scan& z, const LaneIo& e, const Blk& b, int idx /* index() of this sample = pp % 10 */)
{
    int diff = z.si - idx;
    if (diff == 5 || diff == -5) {
        z.csi = clock_dead_reckon(z.ck_sest, z.ck_est, z.ck_count);
        z.si = z.csi;
        return;
    }
    if (idx != z.si) return;
    frame_symbol(z, e, y_sel(b.y, (int) (z.p - b.base)));
}

M17B_FN bool sw_trig(const Scan& z, int k) { return (z.sw >> (8 * k + 4)) & 1; }

// Number of upcoming samples on which the state machine provably does nothing but count, for the
// states the block walk handles (every state but UNLOCKED): `i10` = index() of the next sample.
// advance_quiet() applies the counting.
M17B_FN int quiet_samples(const Scan& z, int i10)
{
    int q;
    switch (z.state) {
    case ST_LSF_SYNC: { int d = z.si - i10; q = d < 0 ? d + 10 : d; break; }       // acts when index() == sample_index
    case ST_STREAM_SYNC: case ST_PACKET_SYNC: case ST_BERT_SYNC:
        q = 77 - z.sync_count; q = q < 0 ? 0 : q;                                  // ++sync_count; acts from 78 on
        break;
    case ST_SYNC_WAIT: q = z.sync_count < 86 ? 86 - z.sync_count : 0; break;       // ++sync_count up to 86, then acts
    case ST_FRAME: {
        int d = z.si - i10; d = d < 0 ? d + 10 : d;                                // samples until index() == sample_index
        int a = d >= 5 ? d - 5 : d + 5;                                            // ... until |sample_index - index()| == 5
        q = d < a ? d : a;
        break;
    }
    default: return 0;
    }
    if (z.need_reset || z.need_update) {                                           // :695-709 acts when index() == 0
        int r = i10 == 0 ? 0 : 10 - i10;
        q = r < q ? r : q;
    }
    return q;
}

M17B_FN void advance_quiet(Scan& z, int n)
{
    z.ck_count += (uint32_t) n;                        // clock_recovery(filtered_sample), ClockRecovery.h:42-45
    if (z.state == ST_STREAM_SYNC || z.state == ST_PACKET_SYNC || z.state == ST_BERT_SYNC || z.state == ST_SYNC_WAIT)
        z.sync_count += n;
    else if (z.state == ST_UNLOCKED && z.missing < 1920)
        z.missing += n;                                // :291-293
}

// n more samples go through the correlator (Correlator.h:43-49: buffer position; the filter's own age)
M17B_FN void clocked(Scan& z, int n)
{
    z.gap_k = z.gap_k + n < kTapsUsed ? z.gap_k + n : kTapsUsed;
    if (n < 10) {
        z.corr_pos += n; if (z.corr_pos >= 80) z.corr_pos -= 80;
        z.i10 += n; if (z.i10 >= 10) z.i10 -= 10;
    } else {
        z.corr_pos = (z.corr_pos + n) % 80;
        z.i10 = z.corr_pos % 10;
    }
}

// M17Demodulator.h:695-712: clock recovery reset / update when index() == 0, then clock_recovery(sample)
M17B_FN void clock_gate(Scan& z, const LaneIo& e, int i10 /* z.pp % 10 */)
{
    uint32_t* sb = e.sb;
    if (i10 == 0) {               // :695-709
        if (z.need_reset) {
            Kalman k; kf_reset(k, (float) z.ssi);          // ClockRecovery.h:33-39
            kf_store(sb, F_CK_X0, k);
            z.ck_count = 0;
            z.csi = z.ssi;
            z.ck_est = 0.f;
            z.need_reset = false;
            z.si = z.ssi;
        } else if (z.need_update) {
            F2 r = clock_update_index(sb, z.ssi, z.ck_count);
            z.ck_sest = r.x;
            z.csi = wrap_index(round_half_away(z.ck_sest));
            z.ck_est = r.y;
            z.ck_count = 0;
            z.need_update = false;
        }
    }
    z.ck_count++;                       // clock_recovery(filtered_sample), :712
}

// One sample of M17Demodulator.h:695-740 (everything between correlator.sample() and the carrier-detect
// cadence check) in the states the block walk handles.
M17B_FN void event_step(Scan& z, const LaneIo& e, const Blk& b, int i10 /* z.pp % 10 */)
{
    clock_gate(z, e, i10);
    switch (z.state) {                  // :714-740
    case ST_UNLOCKED: break;            // never here: the search walk's
    case ST_LSF_SYNC: if (i10 == z.si) do_lsf_sync(z, e, b); break;
    case ST_STREAM_SYNC:
        z.sync_count += 1;
        if (z.sync_count >= 78) do_stream_sync(z, e, b);
        break;
    case ST_PACKET_SYNC:
        z.sync_count += 1;
        if (z.sync_count >= 78) do_packet_or_bert_sync(z, e, b, false);
        break;
    case ST_BERT_SYNC:
        z.sync_count += 1;
        if (z.sync_count >= 78) do_packet_or_bert_sync(z, e, b, true);
        break;
    case ST_SYNC_WAIT:                  // :582-593
        if (z.sync_count < 86) z.sync_count += 1;
        else { z.need_update = true; z.state = ST_FRAME; }
        break;
    default: do_frame(z, e, b, i10); break;
    }
}

struct Chunk { float y[8]; };            // one block of ring_y

M17B_FN void load_chunk(Chunk& c, const LaneIo& e, uint32_t base)
{
    const size_t o = ring_off(base, e.lmask, e.np, e.s);
    const float4* py = reinterpret_cast<const float4*>(e.ring_y + o);
    float4 a0 = py[0], a1 = py[1];
    c.y[0] = a0.x; c.y[1] = a0.y; c.y[2] = a0.z; c.y[3] = a0.w; c.y[4] = a1.x; c.y[5] = a1.y; c.y[6] = a1.z; c.y[7] = a1.w;
}

// the limit filter over block samples [jb, je)
M17B_FN void bq_feed(Biquad& b, const float (&y)[8], int jb, int je)
{
    if (jb == 0 && je == 8) {
#pragma unroll
        for (int i = 0; i != 8; ++i) limit_step(b, y[i]);
    } else {
#pragma unroll
        for (int i = 0; i != 8; ++i)
            if (i >= jb && i < je) limit_step(b, y[i]);
    }
}

// ---------------------------------------------------------------------------
// Carrier-detect pre-pass.  DataCarrierDetect::operator() (DataCarrierDetect.h:53-58) clocks the two-bin
// sliding DFT (SlidingDFT.h:118-132) on EVERY input sample, whatever the demodulator state, and sums
// the bin energies until the next update(), every 384 samples while !dcd_ and every 960 while dcd_
// (M17Demodulator.h:675-689,742-752), after a first window of 1920 + 384 samples.  Every update lands
// on a multiple of 192 samples from the start of the stream, and a window's sums depend only on where
// it starts and how long it is.  dcd_lane runs the recurrence once per submit for all streams in
// lock-step (every lane at the same position: no divergence, none of the scan's register pressure)
// and keeps five staggered sums per bin, one per 192-sample phase of a 960-sample cycle; each sum is
// recorded 384 and 960 samples after its start.  The scan looks the pair up at its update points
// (dcd_sums): the same sequence of float additions as the reference, so the level is bit-identical.
// seg_ring entry q (window starting at sample 192 q): {l1@384, l2@384, l1@960, l2@960}; the first
// window [0, 2304) is stored in entry 0 in place of its unused @384 pair.
// ---------------------------------------------------------------------------
// One lane = one DFT bin of one stream (bin 0 or 1: the two bins and their sums are independent
// recurrences, so a stream is split over two neighbouring lanes that read the same samples).
// The sliding DFT's input x[n] - x[n-120] (SlidingDFT.h:123-124, x = the scaled sample, apps/m17-demod.cpp:489) is
// formed here from the submit's own int16 samples and the raw history tail (`src`: DcdSrc in m17b_kernels.cu, a
// scalar stand-in in tests/emu) -- round 1's matched filter wrote it to a second float ring (4 B/sample of DRAM
// traffic each way and a third of the ring memory).  src.request(base) asks for the 8-sample block at ring position
// `base` (any block that overlaps [head_old, head_old + n_new), or the one after), src.take(dl) delivers it; values
// of samples outside the submit are unspecified and unused.
template <class Src>
M17B_FN void dcd_lane(const DevPtrs& d, uint32_t s, int bin, uint32_t head_old, uint32_t n_new, unsigned long long head64_old, Src& src)
{
    uint32_t* sb = d.st + st_index(s, 0);
    const int f_re = bin ? F_RE1 : F_RE0, f_im = bin ? F_IM1 : F_IM0, f_first = bin ? F_L2 : F_L1, f_acc = bin ? F_ACC2 : F_ACC1;
    float re = u2f(sb[f_re * 32]), im = u2f(sb[f_im * 32]);
    float first = u2f(sb[f_first * 32]);                          // first window [0, 2304)
    float acc[5];
#pragma unroll
    for (int i = 0; i != 5; ++i) acc[i] = u2f(sb[(f_acc + i) * 32]);
    const float cr = M17B_TAB.sdft[2 * bin], ci = M17B_TAB.sdft[2 * bin + 1];
    uint32_t seg = (uint32_t) (head64_old / kSeg);       // ordinal of the segment the next sample is in
    uint32_t age = (uint32_t) (head64_old % kSeg);
    uint32_t m5 = seg % 5u;
    const uint32_t smask = d.n_seg - 1;
    const size_t np = d.np;
    float* const ring = d.seg_ring + (size_t) bin * np + s;       // slot k = bin (@384) / 2 + bin (@960)
    // per sample: SlidingDFT.h:118-132 for this bin, its energy (DataCarrierDetect.h:53-58), the sums
#define M17B_DCD_STEP(x_)                                                        \
    do {                                                                         \
        const float a_ = re + (x_), b_ = im;   /* x_ = sample - oldest */       \
        re = a_ * cr - b_ * ci; im = a_ * ci + b_ * cr;                          \
        const float e_ = re * re + im * im;                                      \
        if (in_first) first += e_;                                               \
        _Pragma("unroll") for (int i_ = 0; i_ != 5; ++i_) acc[i_] += e_;         \
    } while (0)
    // a 192-sample segment ends: record the sums that are 384 / 960 samples old, restart the latter
#define M17B_DCD_BOUNDARY()                                                      \
    do {                                                                         \
        age = 0;                                                                 \
        ++seg;                                  /* the segment that starts with the next sample */ \
        m5 = m5 == 4u ? 0u : m5 + 1u;                                            \
        if (seg == (uint32_t) (kFirstUpdate / kSeg)) { ring[0] = first; in_first = false; } \
        _Pragma("unroll") for (int i_ = 0; i_ != 5; ++i_) {                      \
            const uint32_t ph_ = m5 >= (uint32_t) i_ ? m5 - (uint32_t) i_ : m5 + 5u - (uint32_t) i_; \
            if (ph_ == 2u) ring[(size_t) ((seg - 2u) & smask) * 4 * np] = acc[i_]; \
            if (ph_ == 0u) { ring[((size_t) ((seg - 5u) & smask) * 4 + 2) * np] = acc[i_]; acc[i_] = 0.f; } \
        }                                                                        \
    } while (0)
    bool in_first = seg < (uint32_t) (kFirstUpdate / kSeg);
    const uint32_t n_blk = ((head_old & 7u) + n_new + 7u) >> 3;
    uint32_t base = head_old & ~7u;
    src.request(base);
    for (uint32_t b = 0; b != n_blk; ++b, base += 8u) {
        float dl[8];
        src.take(dl);
        src.request(base + 8u);                             // the next block's samples, before this one's arithmetic
        const uint32_t rel = base - head_old;               // wraps for the partial first block
        if (rel < n_new && n_new - rel >= 8u) {
            // whole block inside the submit; blocks are 8-aligned in stream time, so a segment can only end
            // with the block (192 = 24 blocks)
#pragma unroll
            for (int j = 0; j != 8; ++j) M17B_DCD_STEP(dl[j]);
            age += 8u;
            if (age == (uint32_t) kSeg) M17B_DCD_BOUNDARY();
        } else {
#pragma unroll
            for (int j = 0; j != 8; ++j) {
                if ((uint32_t) (rel + (uint32_t) j) >= n_new) continue;
                M17B_DCD_STEP(dl[j]);
                if (++age == (uint32_t) kSeg) M17B_DCD_BOUNDARY();
            }
        }
    }
#undef M17B_DCD_STEP
#undef M17B_DCD_BOUNDARY
    sb[f_re * 32] = f2u(re); sb[f_im * 32] = f2u(im);
    sb[f_first * 32] = f2u(first);
#pragma unroll
    for (int i = 0; i != 5; ++i) sb[(f_acc + i) * 32] = f2u(acc[i]);
}

// Energy sums of the carrier-detect window of w samples that ends at the scan position `pos`.
M17B_FN void dcd_sums(const DevPtrs& d, const LaneIo& e, uint32_t pos, uint32_t head, unsigned long long head64, int w,
                      float& l1, float& l2)
{
    const unsigned long long u = head64 - (unsigned long long) (head - pos);
    uint32_t q = 0;
    int k = 0;
    if (u != (unsigned long long) kFirstUpdate) {
        q = (uint32_t) ((u - (unsigned long long) w) / kSeg);
        k = w == 384 ? 0 : 2;
    }
    const float* p = d.seg_ring + ((size_t) (q & (d.n_seg - 1)) * 4 + (size_t) k) * d.np + e.s;
    l1 = p[0];
    l2 = p[d.np];
}

// Offset, from the next sample the lane handles (logical slot corr_pos), of the last symbol instant of
// the frame being received (ST_FRAME): symbols are taken when index() == sample_index, 10 samples apart
// (the dead-reckoned sample_index moves by at most +-1 a few times per frame).
constexpr int kWalkBudget = 768;         // samples of block walk per round of an acquiring stream (M17B_STREAMING)
constexpr int kDeferMarginFrame = 6;
constexpr int kDeferMarginSync = 16;
M17B_FN int frame_end_offset(int si, int i10, int framer_index)
{
    int dsi = si - i10;
    if (dsi < 0) dsi += 10;
    return dsi + (183 - (framer_index >> 1)) * 10;
}

// The carrier detect gates the filter and the correlator off after the sample at pos - 1
// (M17Demodulator.h:675-689): freeze what they would still hold when they resume.
M17B_FN_COLD void gate_off(const DevPtrs& d, const LaneIo& e, int gap_k, uint32_t pos)
{
    take_snapshot(e.snap, e.ring_x, e.np, e.lmask, e.s, pos - 1, gap_k);
    // Correlator::buffer_: the last 80 clocked samples, oldest first.  The carrier detect stays on for at
    // least 960 samples (and the start-up clocks 1920), so they are all ring samples.
    if (gap_k < 80) ctr_add(&d.ctr->faults, 1ull);
    for (int j = 0; j != 80; ++j)
        e.csnap[(size_t) j * e.np] = e.ring_y[ring_off(pos - 80u + (uint32_t) j, e.lmask, e.np, e.s)];
}

// First sample in [p, p + n) on which one of the selected sync words triggers according to the search
// pass's masks, as an offset from p; n if there is none.  sel: bits 8 w .. 8 w + 7 set for every wanted word w.
M17B_FN int next_trigger(const LaneIo& e, uint32_t p, int n, uint32_t sel)
{
    int done = 0;
    while (done < n) {
        const uint32_t q = p + (uint32_t) done;
        const uint4 v = *reinterpret_cast<const uint4*>(e.mask_ring + ((size_t) ((q >> 5) & e.lmask) * e.np + e.s) * 4);
        const uint32_t w0 = v.x & sel, w1 = v.y & sel, w2 = v.z & sel, w3 = v.w & sel;
        uint32_t lm = ((w0 | (w0 >> 8) | (w0 >> 16) | (w0 >> 24)) & 0xFFu) | (((w1 | (w1 >> 8) | (w1 >> 16) | (w1 >> 24)) & 0xFFu) << 8) |
                      (((w2 | (w2 >> 8) | (w2 >> 16) | (w2 >> 24)) & 0xFFu) << 16) | (((w3 | (w3 >> 8) | (w3 >> 16) | (w3 >> 24)) & 0xFFu) << 24);
        lm >>= (q & 31u);
        if (lm) {
            const int o = done + first_set(lm);
            return o < n ? o : n;
        }
        done += 32 - (int) (q & 31u);
    }
    return n;
}

// the four trigger bits (bit w = sync word w) the search pass recorded for the sample at p
M17B_FN uint32_t mask_bits(const LaneIo& e, uint32_t p)
{
    const uint32_t m = e.mask_ring[blk_off(p, e.lmask, e.np, e.s)] >> (p & 7u);
    return (m & 1u) | ((m >> 7) & 2u) | ((m >> 14) & 4u) | ((m >> 21) & 8u);
}

// Bring the limit filter from position `from` (where z.bq is valid) to `to` inside a searched span: the
// search pass left the filter history at every block end it covered (ckpt_ring).
M17B_FN void bq_catch_up(Scan& z, const LaneIo& e, uint32_t from, uint32_t to)
{
    if (from == to) return;
    uint32_t start = from;
    const uint32_t tb = to & ~7u;
    if ((int32_t) (tb - from) > 0) {
        const float2 c = e.ckpt_ring[blk_off(tb - 8u, e.lmask, e.np, e.s)];
        z.bq.h1 = c.x; z.bq.h2 = c.y;
        start = tb;
    }
    for (uint32_t p = start; p != to; ++p) limit_step(z.bq, e.ring_y[ring_off(p, e.lmask, e.np, e.s)]);
}

// Diagnostics at the reference's update points (M17Demodulator.h:675-689 while !dcd_: evm argument is
// dev.error() = 0, FreqDevEstimator.h:53; :742-752 while dcd_: evm.evm()), only when the context records them.
M17B_FN_COLD void emit_diag_rec(const DevPtrs& d, uint32_t s, const m17b_diag& v)
{
    const uint32_t slot = ctr_add(&d.ctr->diag_count, 1u);
    if (slot - d.diag_read >= d.diag_cap) { ctr_add(&d.ctr->diag_dropped, 1ull); return; }
    m17b_diag_rec* r = d.diag_recs + (slot & (d.diag_cap - 1));
    r->stream = s; r->pad = 0;
    r->d = v;
}

M17B_FN void prbs_reset(uint32_t* sb)                                 // PRBS9::reset, Util.h:401-411
{
    sb[F_PRBS * 32] = 1u;
    sb[F_BER_BITS * 32] = 0; sb[F_BER_ERRS * 32] = 0;
    for (int i = 0; i != 4; ++i) sb[(F_BER_H0 + i) * 32] = 0;
}

// What apps/m17-demod.cpp's diagnostic callback does besides printing (:353-355): a callback that reports no
// carrier resets the PRBS9 accounting once it has counted anything.  (The callback of a sample runs after that
// sample's frame callback: when the frame it completes is still waiting for decode_kernel, the reset waits too.)
M17B_FN void ber_diag_hook(uint32_t* sb, const Scan& z)
{
    if (z.dcd || sb[F_BER_BITS * 32] == 0) return;
    if (z.pending) sb[F_PRBS * 32] |= 1u << 30;
    else prbs_reset(sb);
}

M17B_FN void emit_diag(const DevPtrs& d, const LaneIo& e, const Scan& z, unsigned long long samples, bool locked_path)
{
    ber_diag_hook(e.sb, z);
    if (!d.diag_recs) return;
    m17b_diag v;
    v.dcd = z.dcd ? 1 : 0;
    v.evm = locked_path ? sqrtf(z.evm) : 0.f;            // SymbolEvm::evm(), StandardDeviation.h:79-82
    v.deviation = 2400.f / z.idev;                       // FreqDevEstimator.h:52
    v.offset = z.offset;
    v.locked = z.state != ST_UNLOCKED;
    v.clock = z.ck_est;
    v.sample_index = z.si; v.sync_index = z.ssi; v.clock_index = z.csi;
    v.viterbi_cost = (int32_t) z.vcost;
    v.samples = samples;
    emit_diag_rec(d, e.s, v);
}

// M17B_STREAMING: would scan_lane leave this stream where it is, because the frame it is in (or waiting for) cannot
// complete before the head?  (classify_kernel asks too, so that such streams take no lane of a round.)
template <typename P>
M17B_FN bool defers_at_once(const P* sb, uint32_t fl, uint32_t pos, uint32_t head)
{
    const int st = (int) ((fl >> FL_STATE_SHIFT) & 7);
    const long long left = (long long) (head - pos);
    if (st == ST_FRAME) {
        const uint32_t idx = sb[F_IDX * 32];
        const int eo = frame_end_offset((int) ((fl >> FL_SI_SHIFT) & 15), (int) (idx & 0xFF) % 10, (int) ((idx >> 8) & 0xFFFF));
        return (long long) eo - kDeferMarginFrame >= left;
    }
    if (st == ST_STREAM_SYNC || st == ST_PACKET_SYNC || st == ST_BERT_SYNC || st == ST_SYNC_WAIT)
        // sync_count counts samples from the last frame's end; frames are 1920 samples apart
        return 1919ll - (long long) sb[F_SYNC_COUNT * 32] - kDeferMarginSync >= left;
    return false;
}

// Walk this lane from its position towards `head`; on return the lane is at head, stopped on a completed
// frame (FL_PENDING, decode_kernel's turn), waiting for fixup_kernel (FL_FIXUP: matched-filter outputs after a
// carrier-detect gap and / or a sync-word search over the samples ahead), or has left a partial frame for
// the next submit (M17B_STREAMING).  Three ways of walking:
//   idle    carrier detect off: nothing is clocked; jump from one 384-sample update to the next;
//   search  ST_UNLOCKED with the carrier detect on: the state machine looks for sync words on every sample.
//           fixup_kernel's search pass evaluates the limit filter and the correlations for the whole span at
//           once (one thread per sample) and leaves trigger masks; the lane jumps from trigger to trigger;
//   block   every other state (and the 1920-sample start-up): aligned 8-sample blocks of ring_y; the limit
//           filter runs densely over the block in registers, the state machine only on the samples it can
//           act on (quiet_samples() gives the distance in closed form).
M17B_FN void scan_lane(const DevPtrs& d, const LaneIo& e, uint32_t head, unsigned long long head64, bool valid, bool defer,
                       uint32_t guard = 0 /* samples a concurrent filter launch may be writing past head */)
{
    if (!valid) return;
    uint32_t* sb = e.sb;
    uint32_t pos = sb[F_POS * 32];
    uint32_t fl = sb[F_FLAGS * 32];
    if ((int32_t) (head - pos) <= 0) return;               // at the head -- or past it: the other chain's rounds may belong to a later submit
    if (fl & (FL_PENDING | FL_FIXUP)) { ctr_add(&d.ctr->cls[e.cls].lagging, 1u); return; }
    // Deferral (rounds enqueued by m17b_submit; m17b_sync's flush rounds run every lane to the head): a lane
    // whose current frame cannot complete before the head leaves the rest of the chunk for the next submit.
    // The records are the same -- every frame that ends inside this chunk is still completed here -- but the
    // lane then walks one whole frame per round together with its neighbours, instead of a partial frame
    // after its last completion and another one at the start of the next chunk.
    // In ST_FRAME the end of the frame is known to +-2 samples (frame_end_offset); right after a frame, while
    // the next sync word is awaited, only to -11 .. +5 (a re-timed sync word moves the symbol phase), hence
    // the wider margin there -- a lane inside that margin starts walking and is stopped by the same test a
    // few blocks later, once it is in ST_FRAME (block walk below).
    if (defer && defers_at_once(sb, fl, pos, head)) { ctr_add(&d.ctr->cls[e.cls].deferred, 1u); return; }
    // a stream that fell so far behind that the sample ring has wrapped over what it still needs
    // (submits enqueued back to back without m17b_sync and more stops than rounds): must never happen
    if (head + guard - pos > d.ring_lines * (uint32_t) kLine - 176u) ctr_add(&d.ctr->faults, 1ull);

    uint32_t gap_r = sb[F_GAPR * 32];                       // position the filter last resumed at
    int gap_done = (int) sb[F_GAPDONE * 32];                // outputs after it already corrected
    uint32_t sfrom = sb[F_SFROM * 32], sto = sb[F_STO * 32];    // span the search pass has covered
    bool fixup = false, deferred = false, yield = false;
    // M17B_STREAMING, acquiring streams: at most kWalkBudget samples of block walk per round.  Waiting for the LSF sync
    // word in a preamble the state machine correlates and updates its estimators on every symbol, an order of magnitude
    // more serial work per sample than inside a frame; a round of the acquiring streams' chain (which has its own rounds,
    // m17b_submit) lasts as long as its slowest walk, so a long one is taken in several rounds (the stream stays `lagging`).
    // A stream that already trails the head by more than a submit walks on regardless (the sample ring is finite).
    int budget = e.split && head - pos <= guard ? kWalkBudget : 0x7fffffff;

    Scan z;
    memset(&z, 0, sizeof(z));
    z.level = u2f(sb[F_LEVEL * 32]);
    z.bq.h1 = u2f(sb[F_IIR0 * 32]); z.bq.h2 = u2f(sb[F_IIR1 * 32]);
    z.idev = u2f(sb[F_IDEV * 32]); z.offset = u2f(sb[F_OFFSET * 32]);
    z.evm = u2f(sb[F_EVM * 32]);
    z.ck_est = u2f(sb[F_CK_EST * 32]); z.ck_sest = u2f(sb[F_CK_SEST * 32]);
    z.ck_count = sb[F_CK_COUNT * 32];
    { uint32_t v = sb[F_CNT * 32]; z.count = v & 0xFFFF; z.init = v >> 16; }
    { uint32_t v = sb[F_IDX * 32]; z.corr_pos = v & 0xFF; z.framer_index = (v >> 8) & 0xFFFF; z.gap_k = v >> 24; }
    z.i10 = z.corr_pos % 10;
    z.state = (fl >> FL_STATE_SHIFT) & 7; z.swt = (fl >> FL_SWT_SHIFT) & 3;
    z.si = (fl >> FL_SI_SHIFT) & 15; z.ssi = (fl >> FL_SSI_SHIFT) & 15; z.csi = (fl >> FL_CSI_SHIFT) & 15;
    z.dcd = fl & FL_DCD; z.dcd_trig = fl & FL_DCD_TRIG; z.need_reset = fl & FL_NEED_RESET;
    z.need_update = fl & FL_NEED_UPDATE; z.dev_reset = fl & FL_DEV_RESET; z.eot = fl & FL_EOT;
    z.sync_count = (int) sb[F_SYNC_COUNT * 32]; z.missing = (int) sb[F_MISSING * 32];
    z.vcost = sb[F_VCOST * 32];
    z.sw = sb[F_SW * 32];

    while (pos != head && !z.pending && !fixup && !deferred && !yield) {
        if (!z.init && !z.dcd) {
            // ---- idle: M17Demodulator.h:675-689 ------------------------------------------------
            const uint32_t left = head - pos;
            const int n = (uint32_t) (384 - z.count) < left ? 384 - z.count : (int) left;
            z.count += n;
            pos += (uint32_t) n;
            if (z.count == 384) {
                float l1, l2;
                dcd_sums(d, e, pos, head, head64, 384, l1, l2);
                update_dcd(z, e);
                dcd_update(z, l1, l2);
                emit_diag(d, e, z, head64 - (unsigned long long) (head - pos), false);
                z.count = 0;
                if (z.dcd) { gap_r = pos; gap_done = 0; sfrom = sto = pos; }   // the filter resumes with the next sample:
                                                            // whatever was searched before the gap no longer applies
            }
            continue;
        }
        if (!z.init && z.state == ST_UNLOCKED) {
            // ---- search: ST_UNLOCKED with the carrier detect on (M17Demodulator.h:288-342) ------
            if (!(pos - sfrom < sto - sfrom)) {             // no masks for this position yet: fixup_kernel's turn
                fixup = true;
                sfrom = sto = pos;
                break;
            }
            const uint32_t left = sto - pos;                // sto <= head
            const int n = (uint32_t) (960 - z.count) < left ? 960 - z.count : (int) left;
            const uint32_t pos0 = pos;
            int j = 0;
            while (j < n) {
                int q;
                if (z.missing < 1920) {
                    if (sw_trig(z, 0)) q = 0;
                    else {
                        const int lim = 1920 - z.missing;   // then the search changes words
                        q = next_trigger(e, pos0 + (uint32_t) j, n - j < lim ? n - j : lim, 0x000000FFu);
                    }
                } else {
                    if (sw_trig(z, 1) || sw_trig(z, 2)) q = 0;
                    else q = next_trigger(e, pos0 + (uint32_t) j, n - j, 0x00FFFF00u);
                }
                if (z.need_reset || z.need_update) {        // :695-709 acts when index() == 0
                    const int r = z.i10 == 0 ? 0 : 10 - z.i10;
                    q = r < q ? r : q;
                }
                if (q >= n - j) { advance_quiet(z, n - j); clocked(z, n - j); j = n; break; }
                if (q) { advance_quiet(z, q); clocked(z, q); j += q; }
                z.p = pos0 + (uint32_t) j;
                z.pp = z.corr_pos;
                const int i10 = z.i10;
                clocked(z, 1);
                z.k = z.gap_k;
                z.mbits = mask_bits(e, z.p);
                z.have_y8 = false;
                clock_gate(z, e, i10);
                do_unlocked(z, e);
                ++j;
                if (z.state != ST_UNLOCKED) break;
            }
            pos = pos0 + (uint32_t) j;
            z.count += j;
            bq_catch_up(z, e, pos0, pos);
            if (z.count == 960) {                           // :742-752
                float l1, l2;
                dcd_sums(d, e, pos, head, head64, 960, l1, l2);
                update_dcd(z, e);
                z.count = 0;
                emit_diag(d, e, z, head64 - (unsigned long long) (head - pos), true);
                dcd_update(z, l1, l2);
                if (!z.dcd) gate_off(d, e, z.gap_k, pos);
            }
            continue;
        }

        if (!z.init && z.state == ST_FRAME && !z.need_reset && !z.need_update && z.gap_k >= kTapsUsed) {
            // ---- frame walk: ST_FRAME, M17Demodulator.h:595-654 ---------------------------------
            // Inside a frame only two kinds of sample do anything: the one half a symbol away from the symbol
            // instant (dead-reckoning clock update, which may move sample_index by one) and the symbol instant
            // itself (slice, soft bits, framer).
            if (defer && head - pos < 1920u &&
                frame_end_offset(z.si, z.i10, z.framer_index) - kDeferMarginFrame >= (int) (head - pos)) {
                deferred = true;                            // this frame ends after the head: the next submit's
                break;
            }
            // M17B_STREAMING: an acquiring stream that has got as far as a frame is a receiving stream from here on:
            // the receiving streams' chain takes it at its next round and walks the frame in step with the others
            if (e.split) { yield = true; break; }
            // Aligned 8-sample blocks, every lane of the warp one block per pass, from here to the end of the
            // frame in ONE loop (the carrier-detect cadence, every 960 samples and always at a block end, is
            // handled inside it), so that the lanes of a warp, which all start their frames in the same pass of
            // the round, stay together: the limit filter over the block, and the block's events -- at most one
            // symbol instant and one clock update in nearly every case, in either order -- in fixed slots
            // (update, symbol, update); anything else falls to the general loop after them.  The symbol's sample
            // comes from the block in registers; blocks are requested two passes ahead.
            uint32_t base = pos & ~7u;
            int jb = (int) (pos - base), done = 0;
            Chunk c0, c1, c2;
            load_chunk(c0, e, base); load_chunk(c1, e, base + 8u); load_chunk(c2, e, base + 16u);
#define M17B_NEXT_EVENTS()                                                                              \
            do { dd = z.si - z.i10; dd = dd < 0 ? dd + 10 : dd;   /* samples until index() == sample_index */  \
                 aa = dd >= 5 ? dd - 5 : dd + 5;                  /* ... until |sample_index - index()| == 5 */ \
            } while (0)
#define M17B_CLOCK_EVENT()                                                                              \
            do { j += aa + 1; z.ck_count += (uint32_t) (aa + 1);  /* clock_recovery(sample) on every sample, :712 */ \
                 z.i10 = z.si >= 4 ? z.si - 4 : z.si + 6;         /* index() of the sample after the one at si + 5 */ \
                 z.si = clock_dead_reckon(z.ck_sest, z.ck_est, z.ck_count); z.csi = z.si;                     \
                 M17B_NEXT_EVENTS();                                                                          \
            } while (0)
#define M17B_SYMBOL_EVENT()                                                                             \
            do { j += dd + 1; z.ck_count += (uint32_t) (dd + 1);                                              \
                 z.i10 = z.si == 9 ? 0 : z.si + 1;                                                            \
                 frame_symbol(z, e, y_sel(c0.y, j - 1));                                                      \
                 M17B_NEXT_EVENTS();                                                                          \
            } while (0)
            for (;;) {
                const uint32_t left = head - pos;
                const int je = left < (uint32_t) (8 - jb) ? jb + (int) left : 8;
                int j = jb, dd, aa;
                M17B_NEXT_EVENTS();
                if (aa < dd && j + aa < je) M17B_CLOCK_EVENT();
                if (dd < aa && j + dd < je) M17B_SYMBOL_EVENT();
                if (!z.pending && aa < dd && j + aa < je) M17B_CLOCK_EVENT();
                while (!z.pending && j + (aa < dd ? aa : dd) < je) {
                    if (aa < dd) M17B_CLOCK_EVENT();
                    else M17B_SYMBOL_EVENT();
                }
                if (!z.pending) {                           // the rest of the block only counts
                    z.ck_count += (uint32_t) (je - j);
                    z.i10 += je - j; z.i10 = z.i10 >= 10 ? z.i10 - 10 : z.i10;
                    j = je;
                }
                bq_feed(z.bq, c0.y, jb, j);
                done += j - jb;
                pos += (uint32_t) (j - jb);
                z.count += j - jb;
                if (z.count == 960) {                       // :742-752
                    float l1, l2;
                    dcd_sums(d, e, pos, head, head64, 960, l1, l2);
                    update_dcd(z, e);
                    z.count = 0;
                    emit_diag(d, e, z, head64 - (unsigned long long) (head - pos), true);
                    dcd_update(z, l1, l2);
                    if (!z.dcd) {
                        z.corr_pos = (z.corr_pos + done) % 80; done = 0;
                        gate_off(d, e, z.gap_k, pos);
                        break;
                    }
                }
                if (z.pending || pos == head || z.state != ST_FRAME) break;
                base += 8u; jb = 0;
                c0 = c1; c1 = c2;
                load_chunk(c2, e, base + 16u);
                if ((base & 31u) == 0u)                     // pull the ring line after the next one into L2
                    prefetch_l2(e.ring_y + ring_off(base + 64u, e.lmask, e.np, e.s));
            }
#undef M17B_NEXT_EVENTS
#undef M17B_CLOCK_EVENT
#undef M17B_SYMBOL_EVENT
            z.corr_pos = (z.corr_pos + done) % 80;
            continue;
        }

        // ---- block walk ------------------------------------------------------------------------
        // pf always holds the block the walk needs next (requested one block of arithmetic earlier)
        Chunk pf;
        load_chunk(pf, e, pos & ~7u);
        for (;;) {
            const uint32_t base = pos & ~7u;
            const int jb = (int) (pos - base);
            const uint32_t left = head - base;
            const int je = left < 8u ? (int) left : 8;
            const bool initializing = z.init != 0;          // :667-673
            if (defer && z.state == ST_FRAME && !initializing && head - pos < 1920u &&
                frame_end_offset(z.si, z.i10, z.framer_index) - kDeferMarginFrame >= (int) (head - pos)) {
                deferred = true;                            // this frame ends after the head: the next submit's
                break;
            }
            if (!initializing && z.gap_k < kTapsUsed) {
                // outputs gap_k .. of the resumed filter are needed: are they corrected yet?
                int k_last = (int) (base + (uint32_t) je - 1u - gap_r);
                if (k_last > kTapsUsed - 1) k_last = kTapsUsed - 1;
                if (k_last >= gap_done) { fixup = true; break; }     // fixup_kernel's turn
            }
            Blk b;
#pragma unroll
            for (int i = 0; i != 8; ++i) b.y[i] = pf.y[i];
            load_chunk(pf, e, base + 8u);                   // the next block's samples, before this block's arithmetic
            b.base = base; b.jb = jb; b.start = z.bq;
            // limit filter over the samples this pass clocks (Correlator.h:43-49)
            if (jb == 0 && je == 8) {
#pragma unroll
                for (int i = 0; i != 8; ++i) limit_step(z.bq, b.y[i]);
            } else {
#pragma unroll
                for (int i = 0; i != 8; ++i)
                    if (i >= jb && i < je) limit_step(z.bq, b.y[i]);
            }
            int j = je;
            if (initializing) {
                const int n = je - jb;                      // count_ stays 0 while initializing
                z.init -= n;
                clocked(z, n);
            } else {
                j = jb;
                while (j < je) {
                    const int q = quiet_samples(z, z.i10);
                    if (q >= je - j) { advance_quiet(z, je - j); clocked(z, je - j); j = je; break; }
                    if (q) { advance_quiet(z, q); clocked(z, q); j += q; }
                    z.p = base + (uint32_t) j;
                    z.pp = z.corr_pos;
                    const int i10 = z.i10;
                    clocked(z, 1);
                    z.k = z.gap_k;
                    z.have_limit = false;
                    z.have_y8 = false;
                    event_step(z, e, b, i10);
                    ++j;
                    if (z.pending || z.state == ST_UNLOCKED) break;
                }
                z.count += j - jb;
                if (j != je) {                              // stopped inside the block: the filter history at that sample
                    z.bq = b.start;
#pragma unroll
                    for (int i = 0; i != 8; ++i)
                        if (i >= jb && i < j) limit_step(z.bq, b.y[i]);
                }
            }
            pos = base + (uint32_t) j;
            if (initializing) {
                if (z.init == 0) {                          // from the next sample on the filter is gated by !dcd_
                    z.gap_k = kTapsUsed;
                    gate_off(d, e, z.gap_k, pos);
                    break;
                }
            } else if (z.count == 960) {                    // :742-752
                float l1, l2;
                dcd_sums(d, e, pos, head, head64, 960, l1, l2);
                update_dcd(z, e);
                z.count = 0;
                emit_diag(d, e, z, head64 - (unsigned long long) (head - pos), true);
                dcd_update(z, l1, l2);
                if (!z.dcd) { gate_off(d, e, z.gap_k, pos); break; }
            }
            if (pos == head || z.pending || z.state == ST_UNLOCKED) break;
            if (!initializing && (budget -= j - jb) <= 0) { yield = true; break; }
            // inside a frame with nothing pending at index() == 0: the rest of the frame is the frame walk's
            if (!initializing && z.state == ST_FRAME && !z.need_reset && !z.need_update && z.gap_k >= kTapsUsed) break;
        }
    }

    // write back
    sb[F_POS * 32] = pos;
    sb[F_LEVEL * 32] = f2u(z.level);
    sb[F_IIR0 * 32] = f2u(z.bq.h1); sb[F_IIR1 * 32] = f2u(z.bq.h2);
    sb[F_IDEV * 32] = f2u(z.idev); sb[F_OFFSET * 32] = f2u(z.offset);
    sb[F_EVM * 32] = f2u(z.evm);
    sb[F_CK_EST * 32] = f2u(z.ck_est); sb[F_CK_SEST * 32] = f2u(z.ck_sest);
    sb[F_CK_COUNT * 32] = z.ck_count;
    sb[F_CNT * 32] = (uint32_t) z.count | ((uint32_t) z.init << 16);
    sb[F_IDX * 32] = (uint32_t) z.corr_pos | ((uint32_t) z.framer_index << 8) | ((uint32_t) z.gap_k << 24);
    fl = ((uint32_t) z.state << FL_STATE_SHIFT) | ((uint32_t) z.swt << FL_SWT_SHIFT) | ((uint32_t) z.si << FL_SI_SHIFT) |
         ((uint32_t) z.ssi << FL_SSI_SHIFT) | ((uint32_t) z.csi << FL_CSI_SHIFT) |
         (z.dcd ? FL_DCD : 0) | (z.dcd_trig ? FL_DCD_TRIG : 0) | (z.need_reset ? FL_NEED_RESET : 0) |
         (z.need_update ? FL_NEED_UPDATE : 0) | (z.dev_reset ? FL_DEV_RESET : 0) | (z.eot ? FL_EOT : 0) |
         (z.pending ? FL_PENDING : 0) | (fixup && e.cls != 0 ? FL_FIXUP : 0);
    sb[F_GAPR * 32] = gap_r;
    sb[F_GAPDONE * 32] = (uint32_t) gap_done;
    sb[F_SYNC_COUNT * 32] = (uint32_t) z.sync_count; sb[F_MISSING * 32] = (uint32_t) z.missing;
    sb[F_SW * 32] = z.sw;
    // the searched span only means something while the search that asked for it goes on
    if (!(z.state == ST_UNLOCKED && z.dcd && !z.init)) sfrom = sto = pos;
    sb[F_SFROM * 32] = sfrom; sb[F_STO * 32] = sto;
    // The flags word goes last, behind a fence: it is what classify_kernel goes by, and a round of the receiving
    // streams may be classified while the acquiring streams' chain of an earlier round is still running (run_round) --
    // whoever sees flags that make the stream its own must see the rest of the state as well.
    publish_fence();
    sb[F_FLAGS * 32] = fl;
    // (a stream that stops for fixup_kernel is no longer `receiving`: if it was walked as one, the acquiring streams'
    // chain picks it up as it stands, stops at once and files the request -- class 0 runs no fix-up kernels)
    if (fixup && e.cls != 0) d.fix_list[(size_t) e.cls * d.np + ctr_add(&d.ctr->cls[e.cls].fix_count, 1u)] = e.s;
    if (z.pending) {
        // queue the frame by what M17FrameDecoder::operator() will do with it (M17FrameDecoder.h:353-392),
        // so that decode_kernel's warps each run one kind
        const int dec_state = (int) (sb[F_DEC * 32] & 0xFF);
        int kind = 4;
        if (z.swt == SW_LSF) kind = 0;
        else if (z.swt == SW_BERT) kind = 3;
        else if (z.swt == SW_STREAM) { if (dec_state == DEC_STREAM) kind = 1; }
        else if (dec_state == DEC_BASIC_PACKET || dec_state == DEC_FULL_PACKET) kind = 2;
        d.dec_list[(size_t) (e.cls * 5 + kind) * d.np + ctr_add(&d.ctr->cls[e.cls].dec_count[kind], 1u)] = e.s;
    }
    if (deferred) ctr_add(&d.ctr->cls[e.cls].deferred, 1u);
    else if (pos != head) ctr_add(&d.ctr->cls[e.cls].lagging, 1u);
}

// Fresh-demodulator state of one lane: the member initialisers of M17Demodulator
// (M17Demodulator.h:148-178) and of its blocks; every other array starts zeroed.
M17B_FN void init_lane(uint32_t* sb, uint32_t pos0 = 0)
{
    for (int f = 0; f != NFIELDS; ++f) sb[f * 32] = 0;
    Kalman k; kf_reset(k, 0.f);                              // KalmanFilter.h:28-31
    kf_store(sb, F_CK_X0, k); kf_store(sb, F_DMIN_X0, k); kf_store(sb, F_DMAX_X0, k);
    sb[F_EVM * 32] = f2u(1.0f);                              // RunningStandardDeviation::S{1.0}
    sb[F_CNT * 32] = 1920u << 16;                            // `initializing`, M17Demodulator.h:659 (per stream)
    sb[F_IDX * 32] = (uint32_t) kTapsUsed << 24;             // no filter gap pending
    sb[F_FLAGS * 32] = FL_DEV_RESET | ((uint32_t) ST_UNLOCKED << FL_STATE_SHIFT) | ((uint32_t) SW_LSF << FL_SWT_SHIFT);
    sb[F_DEC * 32] = DEC_LSF;
    sb[F_PRBS * 32] = 1u;                                    // PRBS9::state{1}
    // ring position of the stream's first sample (0 but for the wrap-around test hook)
    sb[F_POS * 32] = pos0; sb[F_GAPR * 32] = pos0; sb[F_SFROM * 32] = pos0; sb[F_STO * 32] = pos0;
}

M17B_FN LaneIo make_io(const DevPtrs& d, uint32_t s, const LlrTable* tab, int cls = 0, bool split = false)
{
    LaneIo e;
    e.cls = cls;
    e.split = split;
    e.llr_tab = tab;
    e.sb = d.st + st_index(s, 0);
    e.sw_samples = d.sw_samples + s;
    e.llr = d.llr + s;
    e.snap = d.fir_snap + s;
    e.csnap = d.corr_snap + s;
    e.ring_y = d.ring_y;
    e.ring_x = d.ring_x;
    e.mask_ring = d.mask_ring;
    e.ckpt_ring = d.ckpt_ring;
    e.np = d.np;
    e.lmask = d.ring_lines - 1;
    e.s = s;
    return e;
}

// Filter fix-up for one lane (host emulation: all outputs serially; the device kernel in
// m17b_kernels.cu spreads the outputs of one request over a CTA).  Corrects ring_y for the
// outputs [gap_done, min(149, head - R)) after the resume position R.
M17B_FN void fixup_lane(const DevPtrs& d, uint32_t s, uint32_t head)
{
    uint32_t* sb = d.st + st_index(s, 0);
    const uint32_t R = sb[F_GAPR * 32];
    const int done = (int) sb[F_GAPDONE * 32];
    uint32_t avail = head - R;
    const int upto = avail < (uint32_t) kTapsUsed ? (int) avail : kTapsUsed;
    const uint32_t lmask = d.ring_lines - 1;
    if (done >= upto) return;
    float xs[kTapsUsed];
    for (int m = 0; m < upto; ++m) xs[m] = scale_i16(d.ring_x[ring_off(R + (uint32_t) m, lmask, d.np, s)]);
    for (int k = done; k < upto; ++k) {
        float y;
        {
            float result = 0.0f;
            for (int i = 0; i != kTapsUsed; ++i) {
                float v = i <= k ? xs[k - i] : d.fir_snap[(size_t) (i - k - 1) * d.np + s];
                result = add_rn(result, mul_rn(v, M17B_TAB.taps[i]));
            }
            y = result;
        }
        d.ring_y[ring_off(R + (uint32_t) k, lmask, d.np, s)] = y;
    }
    sb[F_GAPDONE * 32] = (uint32_t) upto;
}

// Does the lane that asked for fixup_kernel want a sync-word search over the samples ahead of it?
// (ST_UNLOCKED with the carrier detect on, past the start-up: see scan_lane's search walk.)
M17B_FN bool wants_search(const uint32_t* sb, uint32_t head)
{
    const uint32_t fl = sb[F_FLAGS * 32];
    return ((fl >> FL_STATE_SHIFT) & 7u) == (uint32_t) ST_UNLOCKED && (fl & FL_DCD) && (sb[F_CNT * 32] >> 16) == 0 &&
           sb[F_POS * 32] != head;
}

// End of the span a search request covers: the stream usually finds its sync word within a frame's worth of
// samples, and what lies beyond would be searched for nothing (and keeps the whole round waiting: the limit
// filter over the span is a serial recurrence).  A stream that is still searching at the end asks again.
constexpr uint32_t kSearchSpan = 4096;
M17B_FN uint32_t search_end(uint32_t from, uint32_t head)
{
    return head - from > kSearchSpan ? from + kSearchSpan : head;
}

// One sample of the search pass: the limit filter step and the trigger decisions of the preamble, LSF /
// stream and packet / BERT sync words (the words do_unlocked looks for, M17Demodulator.h:288-342) for the
// sample at p, the k-th clocked one since the last resume.  Returns bit w = word w triggers.
M17B_FN uint32_t search_sample(const LaneIo& e, Biquad& b, uint32_t p, int k)
{
    const float limit = limit_step(b, e.ring_y[ring_off(p, e.lmask, e.np, e.s)]);
    float y8[8];
#pragma unroll
    for (int i = 0; i != 8; ++i) y8[i] = y_back(e, p, k, 70 - 10 * i);
    return trigger_bits(limit, corr8(y8, kWordPreamble), corr8(y8, kWordLsf), corr8(y8, kWordPacket), 0.f);
}

// Search pass for one lane (host emulation: serially; the device kernel spreads it over a CTA): trigger
// masks and limit-filter history at the block ends for [F_POS, head).  Leaves the lane's own state alone.
M17B_FN void search_lane(const DevPtrs& d, uint32_t s, uint32_t head)
{
    uint32_t* sb = d.st + st_index(s, 0);
    if (!wants_search(sb, head)) return;
    const LaneIo e = make_io(d, s, nullptr);
    const uint32_t from = sb[F_POS * 32];
    head = search_end(from, head);
    Biquad b;
    b.h1 = u2f(sb[F_IIR0 * 32]); b.h2 = u2f(sb[F_IIR1 * 32]);
    int k = (int) (sb[F_IDX * 32] >> 24);
    uint32_t word = 0;
    for (uint32_t p = from; p != head; ++p) {
        k = k < kTapsUsed ? k + 1 : k;
        const uint32_t bits = search_sample(e, b, p, k);
        const uint32_t j = p & 7u;
        word |= ((bits & 1u) << j) | (((bits >> 1) & 1u) << (8 + j)) | (((bits >> 2) & 1u) << (16 + j));
        if (j == 7u || p + 1u == head) {
            d.mask_ring[blk_off(p, e.lmask, e.np, s)] = word;
            word = 0;
        }
        if (j == 7u) {
            float2 c; c.x = b.h1; c.y = b.h2;
            d.ckpt_ring[blk_off(p, e.lmask, e.np, s)] = c;
        }
    }
    sb[F_SFROM * 32] = from;
    sb[F_STO * 32] = head;
}

// ---------------------------------------------------------------------------
// frame decode
// ---------------------------------------------------------------------------

constexpr int kDecThreads = 128;
#ifdef __CUDACC__
constexpr int HS = kDecThreads;         // survivor-history column stride
#else
constexpr int HS = 1;
#endif

// CRC16.h:41-62 over the 30 LSF bytes (a column of dec_bytes); 0 means valid.
M17B_FN_COLD uint32_t crc16_lsf(const uint8_t* lsf, size_t np)
{
    uint32_t reg = M17B_TAB.crc_init;
    for (int n = 0; n != 30; ++n) {
        uint32_t byte = lsf[(size_t) n * np];
        for (int i = 0; i != 8; ++i) {
            uint32_t msb = reg & 0x8000;
            reg = ((reg << 1) & 0xFFFF) | ((byte >> (7 - i)) & 1);
            if (msb) reg ^= 0x5935;
        }
    }
    for (int i = 0; i != 16; ++i) {
        uint32_t msb = reg & 0x8000;
        reg = (reg << 1) & 0xFFFF;
        if (msb) reg ^= 0x5935;
    }
    return reg;
}

// One depunctured soft bit through the folded derandomise/deinterleave map, in two halves so that the
// loads of a whole block of trellis steps can be issued before the first one is used.
// branch-free: the load is issued unconditionally (slot 0 for an erasure / the stale element)
M17B_FN int soft_index(int m)
{
    const bool special = m == 0 || m == kMapStale;
    return special ? 0 : (m < 0 ? -m : m) - 1;
}

M17B_FN int load_pair(const int16_t* p)
{
#ifdef __CUDACC__
    // volatile: ptxas otherwise sinks each load to just before its use (register pressure), which exposes
    // a full L2 latency in front of every add-compare-select step (45 % of decode_kernel's stall samples,
    // before r02e, see profiles/README.md); in program order all loads of a block are in flight together
    int v;
    asm volatile("ld.global.s16 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
#else
    return *p;
#endif
}

M17B_FN int soft_finish(int m, int k, int pair, int stale)
{
    int v = (k & 1) ? (int) (int8_t) (pair >> 8) : (int) (int8_t) (pair & 0xFF);
    v = m < 0 ? -v : v;
    v = m == 0 ? 0 : v;
    return m == kMapStale ? stale : v;
}

M17B_FN int soft_bit(const int16_t* llr, size_t np, int m, int stale)
{
    const int k = soft_index(m);
    return soft_finish(m, k, llr[(size_t) (k >> 1) * np], stale);
}

// butterfly J: cost_[J][0] = +7 iff bit 2 of J, cost_[J][1] = +7 iff bit0 ^ bit1 of J
// (Viterbi.h:72-88 with polynomials 031, 027 and memory J << 1).
template <int J>
M17B_FN void butterfly(const int (&m)[16], int (&n)[16], uint32_t& hist, int a0, int a1, int b0, int b1)
{
    constexpr bool c0p = (J >> 2) & 1;
    constexpr bool c1p = ((J ^ (J >> 1)) & 1);
    const int cost0 = (c0p ? a0 : a1) + (c1p ? b0 : b1);     // sum |cost_ - s|
    const int cost1 = (c0p ? a1 : a0) + (c1p ? b1 : b0);     // sum |cost_ + s|
    const int p0 = m[J], p1 = m[J + 8];
    const int m0 = p0 + cost0, m1 = p0 + cost1, m2 = p1 + cost1, m3 = p1 + cost0;
    const bool d0 = m0 > m2, d1 = m1 > m3;                   // Viterbi.h:148-149
    hist |= (uint32_t) d0 << (2 * J);
    hist |= (uint32_t) d1 << (2 * J + 1);
    n[2 * J] = d0 ? m2 : m0;
    n[2 * J + 1] = d1 ? m3 : m1;
}

#ifdef __CUDACC__
// Two states per register: the path metrics fit 16 bits (at most 244 steps x 28, and 16,000 stands in for the
// INT_MAX / 2 of the states a frame cannot start in: such a path still loses to every real one and compares
// with its like exactly as before).  Butterfly J reads states J and J + 8, each broadcast to both halves, adds
// the branch costs (cost0, cost1) / (cost1, cost0) and keeps the smaller: the halves of the result are the new
// states 2J and 2J + 1, and "old path from J + 8 wins" is Viterbi.h:148-149's strict m0 > m2, the complement
// of the a <= b predicate the packed minimum returns.
template <int J>
M17B_FN void butterfly2(const uint32_t (&m)[8], uint32_t (&n)[8], uint32_t& hist, uint32_t c01, uint32_t c10)
{
    const uint32_t p0 = __byte_perm(m[J >> 1], 0, (J & 1) ? 0x3232 : 0x1010);
    const uint32_t p1 = __byte_perm(m[(J + 8) >> 1], 0, (J & 1) ? 0x3232 : 0x1010);
    const uint32_t x = p0 + c01;                               // (m0, m1): halves stay below 2^15, no carry across
    const uint32_t y = p1 + c10;                               // (m2, m3)
    bool le_hi, le_lo;
    n[J] = __vibmin_s16x2(x, y, &le_hi, &le_lo);
    hist |= (le_lo ? 0u : 1u) << (2 * J);
    hist |= (le_hi ? 0u : 1u) << (2 * J + 1);
}
#endif

// Viterbi.h:181-236.  Writes out_bits(kind) bits MSB-first into out[] (to_byte_array,
// Util.h:300-318) and returns round(min_metric / 7.f).
M17B_FN_COLD uint32_t viterbi(int kind, const int16_t* llr, size_t np, int stale,
                              uint16_t* hist_col, int hs /* survivor-history stride */, uint8_t* out)
{
    const int16_t* map = M17B_TAB.vmap[kind];
    const int n_in = map_len(kind), n_out = out_bits(kind), steps = n_in / 2;
#ifdef __CUDACC__
    uint32_t m[8], n[8];                                      // states 2i (low half) and 2i + 1 (high half)
#pragma unroll
    for (int i = 0; i != 8; ++i) m[i] = 16000u | (16000u << 16);
    m[0] = 16000u << 16;
#else
    int m[16], n[16];
#pragma unroll
    for (int i = 0; i != 16; ++i) m[i] = 0x7fffffff / 2;
    m[0] = 0;
#endif
    // Soft bits are fetched a block of 8 trellis steps ahead: 16 independent loads in flight instead of
    // two dependent global loads in front of every add-compare-select step (the kernel is latency-bound:
    // a handful of warps per SM, one frame per thread).
    constexpr int kBlk = 8;
    for (int tb = 0; tb < steps; tb += kBlk) {
        int sbit[2 * kBlk], sm[2 * kBlk];
#pragma unroll
        for (int u = 0; u != 2 * kBlk; ++u) {
            sm[u] = 2 * tb + u < n_in ? map[2 * tb + u] : 0;
            sbit[u] = load_pair(llr + (size_t) (soft_index(sm[u]) >> 1) * np);
        }
#pragma unroll
        for (int u = 0; u != 2 * kBlk; ++u) sbit[u] = soft_finish(sm[u], soft_index(sm[u]), sbit[u], stale);
#pragma unroll
        for (int u = 0; u != kBlk; ++u) {
            const int t = tb + u;
            if (t >= steps) break;
            const int s0 = sbit[2 * u], s1 = sbit[2 * u + 1];
            int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
            if (s0) { a0 = abs(7 - s0); a1 = abs(-7 - s0); }
            if (s1) { b0 = abs(7 - s1); b1 = abs(-7 - s1); }
            uint32_t h = 0;
#ifdef __CUDACC__
            // (cost0, cost1) of a butterfly by its two cost signs (see butterfly<J>): c0p c1p = 00 -> (a1 + b1, a0 + b0),
            // 01 -> (a1 + b0, a0 + b1), 10 and 11 are those with the halves swapped
            const uint32_t pa = (uint32_t) (a1 + b1) | ((uint32_t) (a0 + b0) << 16);
            const uint32_t pb = (uint32_t) (a1 + b0) | ((uint32_t) (a0 + b1) << 16);
            const uint32_t pd = __byte_perm(pa, 0, 0x1032), pc = __byte_perm(pb, 0, 0x1032);
            // butterfly J: c0p = bit 2 of J, c1p = bit0 ^ bit1 of J
            butterfly2<0>(m, n, h, pa, pd); butterfly2<1>(m, n, h, pb, pc);
            butterfly2<2>(m, n, h, pb, pc); butterfly2<3>(m, n, h, pa, pd);
            butterfly2<4>(m, n, h, pc, pb); butterfly2<5>(m, n, h, pd, pa);
            butterfly2<6>(m, n, h, pd, pa); butterfly2<7>(m, n, h, pc, pb);
            hist_col[t * hs] = (uint16_t) h;
#pragma unroll
            for (int i = 0; i != 8; ++i) m[i] = n[i];
#else
            butterfly<0>(m, n, h, a0, a1, b0, b1); butterfly<1>(m, n, h, a0, a1, b0, b1);
            butterfly<2>(m, n, h, a0, a1, b0, b1); butterfly<3>(m, n, h, a0, a1, b0, b1);
            butterfly<4>(m, n, h, a0, a1, b0, b1); butterfly<5>(m, n, h, a0, a1, b0, b1);
            butterfly<6>(m, n, h, a0, a1, b0, b1); butterfly<7>(m, n, h, a0, a1, b0, b1);
            hist_col[t * hs] = (uint16_t) h;
#pragma unroll
            for (int i = 0; i != 16; ++i) m[i] = n[i];
#endif
        }
    }
#ifdef __CUDACC__
    int min_element = 0, min_cost = (int) (m[0] & 0xFFFFu);
#pragma unroll
    for (int i = 0; i != 16; ++i) {
        const int v = (int) ((m[i >> 1] >> (16 * (i & 1))) & 0xFFFFu);
        if (v < min_cost) { min_cost = v; min_element = i; }
    }
#else
    int min_element = 0, min_cost = m[0];
#pragma unroll
    for (int i = 0; i != 16; ++i)
        if (m[i] < min_cost) { min_cost = m[i]; min_element = i; }
#endif
    uint32_t cost = (uint32_t) (long long) roundf((float) min_cost / 7.0f);

    const int n_bytes = (n_out + 7) / 8;
    for (int i = 0; i != n_bytes; ++i) out[i] = 0;
    int next = min_element, index = steps, o = n_out;
    for (int t = steps; t-- != 0 && o != 0;) {
        int v = (hist_col[t * hs] >> next) & 1;      // prevState_[next][v] = (next >> 1) | (v << 3)
        if (index-- <= n_out) { --o; out[o >> 3] |= (uint8_t) ((next & 1) << (7 - (o & 7))); }
        next = (next >> 1) | (v << 3);
    }
    return cost;
}

M17B_FN_COLD void emit_record(const DevPtrs& d, uint32_t s, uint32_t* sb, unsigned long long sample_pos,
                              int type, int cost)
{
    uint32_t seq = sb[F_SEQ * 32];
    sb[F_SEQ * 32] = seq + 1;
    ctr_add(&d.ctr->frames[type], 1ull);
    uint32_t slot = ctr_add(&d.ctr->rec_count, 1u);
    if (slot - d.rec_read >= d.rec_cap) { ctr_add(&d.ctr->dropped, 1ull); return; }
    m17b_frame* r = d.recs + (slot & (d.rec_cap - 1));
    r->stream = s; r->seq = seq; r->sample_pos = sample_pos; r->viterbi_cost = cost;
    r->type = (uint8_t) type; r->pad[0] = r->pad[1] = r->pad[2] = 0;
    // payload[26] + lsf[30] are contiguous in m17b_frame and 4-byte aligned: gather the 56 column bytes
    // with independent loads first, then 14 word stores (byte stores through a uint8_t* may alias the
    // loads, which serialised them: 25 % of decode_kernel's stall samples)
    static_assert(offsetof(m17b_frame, payload) % 4 == 0 && offsetof(m17b_frame, lsf) == offsetof(m17b_frame, payload) + 26 &&
                  sizeof(m17b_frame) == offsetof(m17b_frame, payload) + 56, "record layout");
    const uint8_t* b = d.dec_bytes + s;
    uint32_t wv[14];
#pragma unroll
    for (int i = 0; i != 14; ++i) {
        uint32_t b0 = b[(size_t) (4 * i) * d.np], b1 = b[(size_t) (4 * i + 1) * d.np];
        uint32_t b2 = b[(size_t) (4 * i + 2) * d.np], b3 = b[(size_t) (4 * i + 3) * d.np];
        wv[i] = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
    }
    uint32_t* dst = reinterpret_cast<uint32_t*>(r->payload);
#pragma unroll
    for (int i = 0; i != 14; ++i) dst[i] = wv[i];
}

// unpack_lich, M17FrameDecoder.h:181-215, with Golay24::decode (Golay24.h:203-222).
// Returns false when a codeword is uncorrectable; lich[] holds whatever was
// assembled so far either way, like the reference's output_buffer.lich.
M17B_FN_COLD bool unpack_lich(const int16_t* llr, size_t np, uint8_t* lich)
{
    for (int i = 0; i != 6; ++i) lich[i] = 0;
    int index = 0;
    for (int i = 0; i != 4; ++i) {
        uint32_t codeword = 0;
        for (int j = 0; j != 24; ++j) {
            int v = soft_bit(llr, np, M17B_TAB.lichmap[i * 24 + j], 0);
            codeword <<= 1;
            codeword |= (uint32_t) (v > 0);
        }
        uint32_t synd = (codeword >> 1) & 0xffffffu;          // syndrome(input >> 1), Golay24.h:102-112
        for (int b = 0; b != 12; ++b) { if (synd & 1) synd ^= 0xC75; synd >>= 1; }
        uint32_t decoded = codeword ^ (M17B_GOLAY[synd] << 1);
        if (!(popc(synd) < 3 || !(popc(decoded) & 1))) return false;
        decoded >>= 12;
        if (i & 1) {
            lich[index++] |= (uint8_t) (decoded >> 8);
            lich[index++] = (uint8_t) (decoded & 0xFF);
        } else {
            lich[index++] |= (uint8_t) (decoded >> 4);
            lich[index] = (uint8_t) ((decoded & 0x0F) << 4);
        }
    }
    return true;
}

// ---------------------------------------------------------------------------
// Consumers of the frame records: what apps/m17-demod.cpp keeps per process on top of the decoder's
// callbacks, kept per stream here and fed by decode_lane.
// ---------------------------------------------------------------------------

// dump_lsf's bookkeeping (apps/m17-demod.cpp:154-175): a new LSF clears the packet buffer and the frame
// counter; for a packet-mode LSF (type bit 0 clear) of a packet type other than 1 the buffer starts with the LSF.
// (The reference indexes lsf[109..111] of the 30-byte array there; the intended test is the TYPE field's bits.)
M17B_FN_COLD void consume_lsf(const DevPtrs& d, uint32_t s, uint32_t* sb, const uint8_t* lsf /* column, stride np */, size_t np)
{
    uint8_t* buf = d.pkt_buf + (size_t) s * kPktStride;
    uint32_t len = 0;
    const uint32_t b13 = lsf[(size_t) 13 * np];
    if (!(b13 & 1u) && ((b13 >> 1) & 3u) != 1u) {
        for (int i = 0; i != 30; ++i) buf[i] = lsf[(size_t) i * np];
        len = 30;
    }
    sb[F_PKT * 32] = len | (sb[F_PKT * 32] & 0xFF000000u);           // packet_frame_counter = 0
}

// decode_packet (apps/m17-demod.cpp:207-253): reassembly, sequence check, CRC-16/X.25
// (boost::crc_optimal<16, 0x1021, 0xFFFF, 0xFFFF, true, true>, residue 0x0f47 over data + FCS).
M17B_FN_COLD void consume_packet(const DevPtrs& d, uint32_t s, uint32_t* sb, const uint8_t* seg /* 26 bytes */,
                                 unsigned long long sample_pos)
{
    uint8_t* buf = d.pkt_buf + (size_t) s * kPktStride;
    const uint32_t pk = sb[F_PKT * 32];
    uint32_t len = pk & 0xFFFFu, counter = (pk >> 16) & 0xFFu, ordinal = pk >> 24;
    const uint32_t number = (seg[25] & 0x7Fu) >> 2;
    if (seg[25] & 0x80u) {                                            // last frame of the packet
        const uint32_t n = number < 25u ? number : 25u;
        for (uint32_t i = 0; i != n && len < (uint32_t) M17B_PACKET_MAX; ++i) buf[len++] = seg[i];
        uint32_t reg = 0xFFFFu;
        for (uint32_t i = 0; i != len; ++i) {
            reg ^= buf[i];
            for (int b = 0; b != 8; ++b) reg = (reg & 1u) ? (reg >> 1) ^ 0x8408u : reg >> 1;
        }
        const bool ok = ((~reg) & 0xFFFFu) == 0x0f47u;
        ctr_add(ok ? &d.ctr->packets_ok : &d.ctr->packets_crc_fail, 1ull);
        const uint32_t slot = ctr_add(&d.ctr->pkt_count, 1u);
        if (slot - d.pkt_read >= d.pkt_cap) ctr_add(&d.ctr->packets_dropped, 1ull);
        else {
            m17b_packet* r = d.pkt_recs + (slot & (d.pkt_cap - 1));
            r->stream = s; r->seq = ordinal; r->sample_pos = sample_pos;
            r->length = (uint16_t) len; r->fcs_ok = ok ? 1 : 0; r->pad = 0;
            for (uint32_t i = 0; i != len; ++i) r->data[i] = buf[i];
        }
        ordinal = (ordinal + 1u) & 0xFFu;
    } else if (number != counter) {
        ctr_add(&d.ctr->packet_seq_errors, 1ull);                     // segment dropped, counter unchanged
    } else {
        counter = (counter + 1u) & 0xFFu;
        for (uint32_t i = 0; i != 25u && len < (uint32_t) M17B_PACKET_MAX; ++i) buf[len++] = seg[i];
    }
    sb[F_PKT * 32] = len | (counter << 16) | (ordinal << 24);
}

// PRBS9 (Util.h:320-413) in the per-stream fields; history = 128 bits in F_BER_H0 .. +3
struct Prbs9 {
    uint32_t state, synced, sync_count, hist_count, hist_pos, bits, errs, h[4];
};

// decode_bert (apps/m17-demod.cpp:287-305): 24 bytes + 5 bits through PRBS9::validate
M17B_FN_COLD void consume_bert(uint32_t* sb, const uint8_t* bert /* 25 bytes */)
{
    const uint32_t w = sb[F_PRBS * 32];
    Prbs9 p;
    p.state = w & 0x1FFu; p.synced = (w >> 9) & 1u; p.sync_count = (w >> 10) & 31u;
    p.hist_count = (w >> 15) & 0xFFu; p.hist_pos = (w >> 23) & 0x7Fu;
    p.bits = sb[F_BER_BITS * 32]; p.errs = sb[F_BER_ERRS * 32];
    for (int i = 0; i != 4; ++i) p.h[i] = sb[(F_BER_H0 + i) * 32];
    for (int k = 0; k != 197; ++k) {
        const uint32_t bit = (bert[k >> 3] >> (7 - (k & 7))) & 1u;
        if (!p.synced) {                                              // synchronize(), :365-382
            const uint32_t r = (bit ^ (p.state >> 8) ^ (p.state >> 4)) & 1u;
            p.state = ((p.state << 1) | bit) & 0x1FFu;
            if (r) p.sync_count = 0;
            else if (++p.sync_count == 18u) {
                p.synced = 1; p.bits += 18u;
                p.h[0] = p.h[1] = p.h[2] = p.h[3] = 0;
                p.hist_count = 0; p.hist_pos = 0; p.sync_count = 0;
            }
        } else {                                                      // generate() + count_errors(), :337-361
            const uint32_t g = ((p.state >> 8) ^ (p.state >> 4)) & 1u;
            p.state = ((p.state << 1) | g) & 0x1FFu;
            const uint32_t err = bit ^ g;
            p.bits += 1u;
            const uint32_t wi = p.hist_pos >> 5, m = 1u << (p.hist_pos & 31u);
            p.hist_count -= (p.h[wi] & m) ? 1u : 0u;
            if (err) {
                p.errs += 1u; p.hist_count += 1u; p.h[wi] |= m;
                if (p.hist_count >= 25u) p.synced = 0;
            } else p.h[wi] &= ~m;
            if (++p.hist_pos == 128u) p.hist_pos = 0;
        }
    }
    if ((w >> 30) & 1u) { prbs_reset(sb); return; }                   // the diagnostic callback of this very sample saw no carrier
    sb[F_PRBS * 32] = p.state | (p.synced << 9) | (p.sync_count << 10) | ((p.hist_count & 0xFFu) << 15) | (p.hist_pos << 23);
    sb[F_BER_BITS * 32] = p.bits; sb[F_BER_ERRS * 32] = p.errs;
    for (int i = 0; i != 4; ++i) sb[(F_BER_H0 + i) * 32] = p.h[i];
}

// M17FrameDecoder::operator() for one lane whose framer has completed a frame.
M17B_FN void decode_lane(const DevPtrs& d, uint32_t s, uint16_t* hcol, uint32_t head, unsigned long long head64, int hs = HS)
{
    uint32_t* sb = d.st + st_index(s, 0);
    uint32_t fl = sb[F_FLAGS * 32];
    if (!(fl & FL_PENDING)) return;

    const size_t np = d.np;
    const int16_t* llr = d.llr + s;
    uint8_t* bytes = d.dec_bytes + s;                      // rows 0..25 union payload, 26..55 lsf
    const uint32_t dec = sb[F_DEC * 32];
    int dec_state = dec & 0xFF;
    uint32_t lich_segments = (dec >> 8) & 0xFF;
    int stale = (int) (int8_t) ((dec >> 16) & 0xFF);       // depuncture_buffer element 401
    uint32_t vcost = sb[F_VCOST * 32];
    const int swt = (fl >> FL_SWT_SHIFT) & 3;
    // the frame completed on the last sample the scan consumed: position pos - 1
    const unsigned long long sample_pos = head64 - (unsigned long long) (head - sb[F_POS * 32]) - 1ull;

    uint8_t out[30];

    // M17FrameDecoder.h:353-392
    if (swt == SW_LSF) {
        dec_state = DEC_LSF;
        // decode_lsf, :154-178
        vcost = viterbi(0, llr, np, stale, hcol, hs, out);
        stale = soft_bit(llr, np, M17B_TAB.vmap[0][401], 0);              // depuncture_buffer.lsf[401]
        for (int i = 0; i != 30; ++i) bytes[(size_t) (26 + i) * np] = out[i];
        if (crc16_lsf(bytes + 26 * np, np) == 0) {
            // update_state, :113-136 (bit 111 = byte 13 bit 0, 109 = bit 2, 110 = bit 1)
            int b111 = out[13] & 1, b109 = (out[13] >> 2) & 1, b110 = (out[13] >> 1) & 1;
            if (b111) { if (b109) dec_state = DEC_STREAM; }
            else dec_state = ((b109 << 1) | b110) == 1 ? DEC_BASIC_PACKET : DEC_FULL_PACKET;
            emit_record(d, s, sb, sample_pos, M17B_FRAME_LSF, (int) vcost);
            consume_lsf(d, s, sb, bytes + 26 * np, np);
        } else {
            ctr_add(&d.ctr->lsf_crc_fail, 1ull);
            lich_segments = 0;
            for (int i = 0; i != 30; ++i) bytes[(size_t) (26 + i) * np] = 0;
        }
    } else if (swt == SW_STREAM) {
        if (dec_state == DEC_LSF) {
            // decode_lich, :217-262
            uint8_t lich[6];
            bool ok = unpack_lich(llr, np, lich);
            for (int i = 0; i != 6; ++i) bytes[(size_t) i * np] = lich[i];
            if (!ok) {
                ctr_add(&d.ctr->golay_fail, 1ull);
            } else {
                emit_record(d, s, sb, sample_pos, M17B_FRAME_LICH, 0);
                uint32_t fragment = (lich[5] >> 5) & 7;
                if (fragment > 5) {
                    vcost = 0xFFFFFFFFu;
                } else {
                    for (int i = 0; i != 5; ++i) bytes[(size_t) (26 + fragment * 5 + i) * np] = lich[i];
                    lich_segments |= 1u << fragment;
                    if ((lich_segments & 0x3F) != 0x3F) {
                        vcost = 0xFFFFFFFFu;
                    } else if (crc16_lsf(bytes + 26 * np, np) == 0) {
                        lich_segments = 0;
                        dec_state = DEC_STREAM;
                        vcost = 0;
                        emit_record(d, s, sb, sample_pos, M17B_FRAME_LSF, 0);
                        consume_lsf(d, s, sb, bytes + 26 * np, np);
                    } else {
                        vcost = 128;
                    }
                }
            }
        } else if (dec_state == DEC_STREAM) {
            // decode_stream, :276-289
            vcost = viterbi(1, llr, np, stale, hcol, hs, out);
            for (int i = 0; i != 18; ++i) bytes[(size_t) i * np] = out[i];
            emit_record(d, s, sb, sample_pos, M17B_FRAME_STREAM, (int) vcost);
        } else {
            dec_state = DEC_LSF;
        }
    } else if (swt == SW_PACKET) {
        if (dec_state == DEC_BASIC_PACKET || dec_state == DEC_FULL_PACKET) {
            // decode_packet, :299-315
            vcost = viterbi(2, llr, np, stale, hcol, hs, out);
            stale = soft_bit(llr, np, M17B_TAB.vmap[2][401], 0);          // depuncture_buffer.packet[401]
            for (int i = 0; i != 26; ++i) bytes[(size_t) i * np] = out[i];
            emit_record(d, s, sb, sample_pos,
                        dec_state == DEC_BASIC_PACKET ? M17B_FRAME_BASIC_PACKET : M17B_FRAME_FULL_PACKET, (int) vcost);
            consume_packet(d, s, sb, out, sample_pos);
            if (out[25] & 0x80) dec_state = DEC_LSF;
        } else {
            dec_state = DEC_LSF;
        }
    } else {
        // decode_bert, :264-274
        dec_state = DEC_BERT;
        vcost = viterbi(3, llr, np, stale, hcol, hs, out);
        for (int i = 0; i != 25; ++i) bytes[(size_t) i * np] = out[i];
        emit_record(d, s, sb, sample_pos, M17B_FRAME_BERT, (int) vcost);
        consume_bert(sb, out);
    }
    if (swt != SW_BERT && (sb[F_PRBS * 32] >> 30) & 1u) prbs_reset(sb);    // (reset deferred behind this frame, see ber_diag_hook)

    sb[F_DEC * 32] = (uint32_t) dec_state | (lich_segments << 8) | ((uint32_t) (uint8_t) (int8_t) stale << 16);
    sb[F_VCOST * 32] = vcost;
    // M17Demodulator.h:627-642, unless dcd_off already forced UNLOCKED on the same sample
    int state = (fl >> FL_STATE_SHIFT) & 7;
    if (state == ST_TBD) {
        switch (dec_state) {
        case DEC_STREAM: state = ST_STREAM_SYNC; break;
        case DEC_LSF: state = ST_STREAM_SYNC; break;
        case DEC_BERT: state = ST_BERT_SYNC; break;
        default: state = ST_PACKET_SYNC; break;
        }
    }
    fl = (fl & ~(7u << FL_STATE_SHIFT) & ~FL_PENDING) | ((uint32_t) state << FL_STATE_SHIFT);
    publish_fence();                                        // see scan_lane's write-back
    sb[F_FLAGS * 32] = fl;
}

} // namespace m17b
