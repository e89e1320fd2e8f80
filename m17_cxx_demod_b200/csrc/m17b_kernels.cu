// m17b_kernels.cu -- hand-written sm_100a kernels of the M17 demodulation path.
//
//   fir_kernel     (A) int16 -> float scaling (apps/m17-demod.cpp:489) + 150-tap RRC
//                      matched filter (FirFilter.h:28-43, taps M17Demodulator.h:79-118)
//                      for every stream and every new sample; FP32 pipe.
//   scan_kernel    (B) per-stream serial part of M17Demodulator::operator()
//                      (M17Demodulator.h:656-753): sliding-DFT carrier detect, limit
//                      biquad, sync-word correlators, Kalman clock recovery, deviation /
//                      offset estimator, 4-FSK soft slicer, framer.  One thread per
//                      stream, every lane at its own stream position.
//   decode_kernel  (C) M17FrameDecoder::operator() (M17FrameDecoder.h:353-392):
//                      derandomise + deinterleave + depuncture folded into one index
//                      map, soft Viterbi, Golay24 LICH, CRC16, decoder state machine,
//                      frame-record emission.
//
// Compiled with -fmad=false: every float operation below is a separately rounded
// IEEE operation unless it is an explicit __fmaf_rn.
#include "m17b_dev.cuh"

#include <cstdio>

namespace m17b {

// ---------------------------------------------------------------------------
// constants
// ---------------------------------------------------------------------------

__constant__ ConstTables c_tab;
__device__ uint32_t g_golay[2048];

// Taps as compile-time literals so the fully unrolled filter loop carries them as
// instruction immediates (FFMA/FMUL imm form).  Same values as ConstTables::taps.
#define M17B_RRC_HALF \
    0.0029364388513841593, 0.0031468394550958484, 0.002699564567597445, 0.001661182944400927, \
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

__host__ __device__ constexpr float rrc_tap(int i)
{
    constexpr double half[75] = {M17B_RRC_HALF};
    return i >= 149 ? 0.0f : float(i <= 74 ? half[i] : half[148 - i]);
}

cudaError_t upload_tables(const ConstTables& t, const uint32_t* golay)
{
    cudaError_t e = cudaMemcpyToSymbol(c_tab, &t, sizeof(t));
    if (e != cudaSuccess) return e;
    return cudaMemcpyToSymbol(g_golay, golay, 2048 * sizeof(uint32_t));
}

// ---------------------------------------------------------------------------
// (A) scaling + matched filter
// ---------------------------------------------------------------------------

// float(double(s) / 41067.0) without double arithmetic: q = s*r, one Newton
// correction with two FMAs gives the correctly rounded float quotient, which
// equals the reference's double-rounded value for all 65536 inputs
// (checked exhaustively in tests/test_scale.py and on the device in the GPU tests).
__device__ __forceinline__ float scale_sample(int v)
{
    const float r = 0x1.988864p-16f;       // RN(1 / 41067)
    float sf = (float) v;
    float q = __fmul_rn(sf, r);
    float e = __fmaf_rn(-q, 41067.0f, sf);
    return __fmaf_rn(e, r, q);
}

constexpr int kFirStreams = 32;     // streams per CTA (one per lane)
constexpr int kFirWarps = 8;
constexpr int kFirOut = 16;         // outputs per thread
constexpr int kFirTile = kFirWarps * kFirOut;      // 128 positions per CTA
constexpr int kFirHalo = 160;
constexpr int kFirRow = kFirTile + kFirHalo + 1;   // 289: odd stride -> conflict-free columns

template <int I, bool FMA>
struct FirStep {
    // Tap I of the filter for 16 consecutive outputs held in acc[]; w[] is a
    // 16-entry register window: slot (k & 15) holds sample t0 + k, k = r - I.
    __device__ __forceinline__ static void run(float (&acc)[kFirOut], float (&w)[kFirOut], const float* row)
    {
        if constexpr (I > 0) w[(16 - (I & 15)) & 15] = row[-I];
        constexpr float tap = rrc_tap(I);
#pragma unroll
        for (int r = 0; r != kFirOut; ++r) {
            float x = w[(r - I) & 15];
            acc[r] = FMA ? __fmaf_rn(x, tap, acc[r]) : __fadd_rn(acc[r], __fmul_rn(x, tap));
        }
        if constexpr (I + 1 < kTapsUsed) FirStep<I + 1, FMA>::run(acc, w, row);
    }
};

template <bool FMA>
__global__ void __launch_bounds__(kFirWarps * 32)
fir_kernel(DevPtrs d, const int16_t* __restrict__ in, size_t stride, uint32_t n_new,
           const int16_t* __restrict__ tail, uint32_t head, int invert)
{
    extern __shared__ float xs[];                  // [32][kFirRow]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t s0 = blockIdx.x * kFirStreams;
    const uint32_t b0 = head & ~7u;
    const uint32_t tb = b0 + blockIdx.y * kFirTile;          // first position of this tile
    const int rel0 = (int) (tb - head) - kFirHalo;            // input index of tile column 0

    // stage [tb-160, tb+128) of 32 streams, scaled to float
    for (int e = tid; e < kFirStreams * (kFirTile + kFirHalo); e += kFirWarps * 32) {
        int sl = e / (kFirTile + kFirHalo), t = e - sl * (kFirTile + kFirHalo);
        uint32_t s = s0 + sl;
        int rel = rel0 + t;
        int v = 0;
        if (s < d.n_streams) {
            if (rel >= 0) { if ((uint32_t) rel < n_new) v = in[(size_t) s * stride + rel]; }
            else if (rel >= -kTailLen) v = tail[(size_t) s * kTailLen + (kTailLen + rel)];
        }
        if (invert) v = (int) (int16_t) (-v);                 // int16 `sample *= -1`
        xs[sl * kFirRow + t] = scale_sample(v);
    }
    __syncthreads();

    const float* row = xs + lane * kFirRow + kFirHalo + warp * kFirOut;     // sample t0
    float acc[kFirOut], w[kFirOut];
#pragma unroll
    for (int r = 0; r != kFirOut; ++r) { acc[r] = 0.0f; w[r] = row[r]; }
    FirStep<0, FMA>::run(acc, w, row);

    const uint32_t s = s0 + lane;
    const uint32_t c = ((tb >> 3) + warp * (kFirOut / 8)) & (d.ring_chunks - 1);
#pragma unroll
    for (int k = 0; k != kFirOut / 8; ++k) {
        uint32_t cc = (c + k) & (d.ring_chunks - 1);
        size_t o = ((size_t) cc * d.np + s) * 8;
        float4* py = reinterpret_cast<float4*>(d.ring_y + o);
        float4* px = reinterpret_cast<float4*>(d.ring_x + o);
        py[0] = make_float4(acc[8 * k + 0], acc[8 * k + 1], acc[8 * k + 2], acc[8 * k + 3]);
        py[1] = make_float4(acc[8 * k + 4], acc[8 * k + 5], acc[8 * k + 6], acc[8 * k + 7]);
        px[0] = make_float4(row[8 * k + 0], row[8 * k + 1], row[8 * k + 2], row[8 * k + 3]);
        px[1] = make_float4(row[8 * k + 4], row[8 * k + 5], row[8 * k + 6], row[8 * k + 7]);
    }
}

void launch_fir(const DevPtrs& d, const int16_t* in, size_t stride, uint32_t n_new, const int16_t* tail,
                uint32_t head, bool fma, bool invert, cudaStream_t s)
{
    uint32_t b0 = head & ~7u;
    uint32_t span = head + n_new - b0;
    dim3 grid(d.np / kFirStreams, (span + kFirTile - 1) / kFirTile);
    size_t smem = sizeof(float) * kFirStreams * kFirRow;
    if (fma) fir_kernel<true><<<grid, kFirWarps * 32, smem, s>>>(d, in, stride, n_new, tail, head, invert);
    else fir_kernel<false><<<grid, kFirWarps * 32, smem, s>>>(d, in, stride, n_new, tail, head, invert);
}

// Keep the last kTailLen raw samples of [tail | new chunk] per stream.
__global__ void tail_kernel(uint32_t n_streams, const int16_t* __restrict__ in, size_t stride, uint32_t n_new,
                            const int16_t* __restrict__ tail_in, int16_t* __restrict__ tail_out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_streams * kTailLen) return;
    uint32_t s = i / kTailLen, k = i - s * kTailLen;
    int rel = (int) n_new - kTailLen + (int) k;        // index into the new chunk
    int16_t v;
    if (rel >= 0) v = in[(size_t) s * stride + rel];
    else v = tail_in[(size_t) s * kTailLen + (kTailLen + rel)];     // rel >= -kTailLen always
    tail_out[i] = v;
}

void launch_tail(const DevPtrs& d, const int16_t* in, size_t stride, uint32_t n_new, const int16_t* tail_in,
                 int16_t* tail_out, cudaStream_t s)
{
    uint32_t n = d.n_streams * kTailLen;
    tail_kernel<<<(n + 255) / 256, 256, 0, s>>>(d.n_streams, in, stride, n_new, tail_in, tail_out);
}

// ring -> [stream][n] for m17b_debug_filter
__global__ void gather_ring_kernel(DevPtrs d, uint32_t n, float* __restrict__ y, float* __restrict__ x)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t s = blockIdx.y;
    if (i >= n || s >= d.n_streams) return;
    size_t o = ((size_t) ((i >> 3) & (d.ring_chunks - 1)) * d.np + s) * 8 + (i & 7);
    y[(size_t) s * n + i] = d.ring_y[o];
    if (x) x[(size_t) s * n + i] = d.ring_x[o];
}

void launch_gather(const DevPtrs& d, uint32_t n, float* y, float* x, cudaStream_t s)
{
    dim3 grid((n + 255) / 256, d.n_streams);
    gather_ring_kernel<<<grid, 256, 0, s>>>(d, n, y, x);
}

// ---------------------------------------------------------------------------
// (B) per-stream scan
// ---------------------------------------------------------------------------

constexpr int kScanThreads = 128;

struct Kalman { float x0, x1, p00, p01, p10, p11; };

__device__ __forceinline__ Kalman kf_load(const uint32_t* sb, int f)
{
    Kalman k;
    k.x0 = __uint_as_float(sb[(f + 0) * 32]); k.x1 = __uint_as_float(sb[(f + 1) * 32]);
    k.p00 = __uint_as_float(sb[(f + 2) * 32]); k.p01 = __uint_as_float(sb[(f + 3) * 32]);
    k.p10 = __uint_as_float(sb[(f + 4) * 32]); k.p11 = __uint_as_float(sb[(f + 5) * 32]);
    return k;
}

__device__ __forceinline__ void kf_store(uint32_t* sb, int f, const Kalman& k)
{
    sb[(f + 0) * 32] = __float_as_uint(k.x0); sb[(f + 1) * 32] = __float_as_uint(k.x1);
    sb[(f + 2) * 32] = __float_as_uint(k.p00); sb[(f + 3) * 32] = __float_as_uint(k.p01);
    sb[(f + 4) * 32] = __float_as_uint(k.p10); sb[(f + 5) * 32] = __float_as_uint(k.p11);
}

// KalmanFilter.h:33-38 / :83-88
__device__ __forceinline__ void kf_reset(Kalman& k, float z)
{
    k.x0 = z; k.x1 = 0.f; k.p00 = 4.f; k.p01 = 0.f; k.p10 = 0.f; k.p11 = 0.00000025f;
}

// KalmanFilter.h:40-65 (wrap) / :90-107.  The reference evaluates this through
// blaze expression templates; operation order and precision below are those of
// the compiled reference (float 2-term sums; the gain scaled in double and
// rounded to float before the state add; the covariance update in double).
// The multiplications by H = [1 0] and F = [[1 dt][0 1]] entries are kept so
// that NaN/Inf propagate identically.
__device__ __forceinline__ void kf_update(Kalman& k, float z, float dt, bool wrap)
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

__device__ __forceinline__ int wrap_index(int v)      // ClockRecovery.h:61-62 on int8
{
    v = v < 0 ? v + 10 : v;
    v = v >= 10 ? v - 10 : v;
    return v;
}

// ---- cold paths: by-value arguments so the hot state stays in registers -----

// Correlator.h:51-64.  ring = this thread's smem column (stride kScanThreads).
__device__ __forceinline__ float correlate(const float* ring, int prev_pos, uint32_t word /* bit i: sync[i] = -3 */)
{
    float result = 0.f;
    int pos = prev_pos + 10;
#pragma unroll
    for (int i = 0; i != 8; ++i) {
        if (pos >= 80) pos -= 80;
        float sy = ((word >> i) & 1) ? -3.f : 3.f;
        result = result + sy * ring[pos * kScanThreads];
        pos += 10;
    }
    return result;
}

// sync words, bit i set <=> symbol i is -3 (M17Demodulator.h:154-157)
constexpr uint32_t kWordPreamble = 0xAA;   // +3,-3,+3,-3,+3,-3,+3,-3
constexpr uint32_t kWordLsf = 0xB0;        // +3,+3,+3,+3,-3,-3,+3,-3
constexpr uint32_t kWordPacket = 0xF2;     // +3,-3,+3,+3,-3,-3,-3,-3
constexpr uint32_t kWordEot = 0x40;        // +3,+3,+3,+3,+3,+3,-3,+3
constexpr float kLowest = -3.402823466e+38f;   // numeric_limits<float>::lowest(), Correlator.h:143

// Correlator.h:150-157
__device__ __forceinline__ float sync_triggered(const float* ring, int prev_pos, float limit, uint32_t word,
                                                float m1, float m2)
{
    float limit_1 = limit * m1;
    float limit_2 = limit * m2;
    float value = correlate(ring, prev_pos, word);
    return (value > limit_1 || value < limit_2) ? value : 0.0f;
}

// Correlator.h:161-207: SyncWord::operator() followed by updated().  `f` is the
// packed {timing_index:4, triggered:1}; returns new f | (updated & 3) << 8
// (updated: 0, 1, or 3 for -1).  samples = SyncWord::samples_ column, stride np.
__device__ __noinline__ uint32_t sync_step(float* samples, size_t np, const float* ring, int prev_pos, float limit,
                                           uint32_t word, float m1, float m2, uint32_t f)
{
    float value = sync_triggered(ring, prev_pos, limit, word, m1, m2);
    uint32_t timing_index = f & 0xF;
    bool trig = (f >> 4) & 1;
    uint32_t updated = 0;
    if (value != 0) {
        if (!trig) {
            for (int i = 0; i != 10; ++i) samples[(size_t) i * np] = 0.f;
            trig = true;
        }
        samples[(size_t) (prev_pos % 10) * np] = value;
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
// Returns (offset, idev); dev.reset_ is always false afterwards.
__device__ __noinline__ float2 update_values(uint32_t* sb, const float* ring, int si, bool dev_reset)
{
    float min_sum = 0, max_sum = 0;
    int min_count = 0, max_count = 0;
    float min_level = ring[si * kScanThreads], max_level = min_level;
    for (int i = si; i < 80; i += 10) {
        float v = ring[i * kScanThreads];
        min_level = v < min_level ? v : min_level;
        max_level = max_level < v ? v : max_level;
    }
    float avg = (float) ((double) max_level + (double) min_level / 2.);
    for (int i = si; i < 80; i += 10) {
        float v = ring[i * kScanThreads];
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
    return make_float2(offset, idev);
}

// ClockRecovery.h:54-67: Kalman update with the sync word's index; returns
// (sample_estimate_, clock_estimate_).
__device__ __noinline__ float2 clock_update_index(uint32_t* sb, int ssi, uint32_t ck_count)
{
    Kalman k = kf_load(sb, F_CK_X0);
    kf_update(k, (float) ssi, (float) ck_count, true);
    kf_store(sb, F_CK_X0, k);
    return make_float2(k.x0, k.x1);
}

__device__ __forceinline__ size_t ring_off(uint32_t p, uint32_t cmask, uint32_t np, uint32_t s)
{
    return ((size_t) ((p >> 3) & cmask) * np + s) * 8 + (p & 7);
}

// FIR history snapshot when the carrier detect gates the filter off
// (M17Demodulator.h:675-689: demod_filter is not clocked while !dcd_).  snap[j] =
// the j-th most recent input the filter has seen; k = samples clocked since the
// last resume (>= 149 unless the carrier dropped again at once).
__device__ __noinline__ void take_snapshot(float* snap, const float* ring_x, uint32_t np, uint32_t cmask,
                                           uint32_t s, uint32_t pos, int k)
{
    if (k < kTapsUsed)
        for (int j = kTapsUsed - 1; j >= k; --j) snap[(size_t) j * np] = snap[(size_t) (j - k) * np];
    int n = k < kTapsUsed ? k : kTapsUsed;
    for (int j = 0; j != n; ++j) snap[(size_t) j * np] = ring_x[ring_off(pos - (uint32_t) j, cmask, np, s)];
}

// Exact matched filter for the first 149 samples after the filter resumes: taps
// reach back across the gap into the snapshot (FirFilter.h:28-43 on stale history).
__device__ __noinline__ float slow_fir(const float* snap, const float* ring_x, uint32_t np, uint32_t cmask,
                                       uint32_t s, uint32_t pos, int k)
{
    float result = 0.0f;
    for (int i = 0; i != kTapsUsed; ++i) {
        float v = i <= k ? ring_x[ring_off(pos - (uint32_t) i, cmask, np, s)] : snap[(size_t) (i - k - 1) * np];
        result = __fadd_rn(result, __fmul_rn(v, c_tab.taps[i]));
    }
    return result;
}

// Util.h:128-145 (llr<float,4>) -> packed (second << 8 | first) int8 pair.  The reference
// binary-searches 43 thresholds spaced ~1/7 apart; the same lower_bound index is found here
// from an arithmetic first guess corrected against the thresholds themselves (a copy of the
// table in shared memory: lanes look up different entries).
struct LlrTable { float thr[44]; uint32_t val[44]; };

__device__ __forceinline__ uint32_t llr4(const LlrTable& tab, float sample)
{
    float t = (-3.0f < sample) ? sample : -3.0f;      // std::max(MIN_VALUE, sample)
    float v = (t < 3.0f) ? t : 3.0f;                  // std::min(MAX_VALUE, t)
    int g = (int) ((v + 3.0f) * 7.0f);                // about the number of thresholds below v
    g = g < 0 ? 0 : (g > 43 ? 43 : g);
    while (g > 0 && !(tab.thr[g - 1] < v)) --g;       // lower_bound: first index with !(thr < v)
    while (g < 43 && tab.thr[g] < v) ++g;
    if (g == 43) g = 42;
    return tab.val[g];
}

// Per-thread working set of the scan; only ever touched by inlined code.
struct Scan {
    float re0, im0, re1, im1, l1, l2, level;   // NSlidingDFT bins, DCD accumulators, DCD level
    float iir0, iir1, limit;                   // limit biquad history / output
    float idev, offset, evm;
    float ck_est, ck_sest;
    uint32_t ck_count;
    int count, init;
    int corr_pos, prev_pos, framer_index, gap_k;
    int state, swt, si, ssi, csi;
    bool dcd, dcd_trig, need_reset, need_update, dev_reset, eot, pending;
    int sync_count, missing;
    uint32_t vcost;
    uint32_t sw;                               // packed SyncWord state, 8 bits per word
};

constexpr size_t kScanSmem = sizeof(float) * (80 + 24) * kScanThreads + sizeof(LlrTable);

struct ScanEnv {
    const LlrTable* llr_tab;
    uint32_t* sb;          // &st[block][0][lane]
    float* ring;           // smem column
    float* sw_samples;     // + s
    int16_t* llr;          // + s
    uint32_t np;
};

#define M17B_UPDATE_VALUES(index_)                                        \
    do {                                                                  \
        float2 od_ = update_values(e.sb, e.ring, z.si, z.dev_reset);      \
        z.offset = od_.x; z.idev = od_.y; z.dev_reset = false;            \
        z.ssi = (index_);                                                 \
    } while (0)

__device__ __forceinline__ int do_sync_step(Scan& z, const ScanEnv& e, int k, uint32_t word, float m1, float m2,
                                            int* timing)
{
    uint32_t f = (z.sw >> (8 * k)) & 0xFF;
    uint32_t r = sync_step(e.sw_samples + (size_t) (10 * k) * e.np, e.np, e.ring, z.prev_pos, z.limit, word, m1, m2, f);
    z.sw = (z.sw & ~(0xFFu << (8 * k))) | ((r & 0xFF) << (8 * k));
    *timing = (int) (r & 0xF);
    uint32_t u = (r >> 8) & 3;
    return u == 0 ? 0 : (u == 1 ? 1 : -1);
}

// DataCarrierDetect.h:63-69 (ltrigger 0.1, htrigger 4.0)
__device__ __forceinline__ void dcd_update(Scan& z)
{
    z.level = (float) ((double) z.level * 0.8 + 0.2 * (double) (z.l1 / z.l2));
    z.l1 = 0.f; z.l2 = 0.f;
    z.dcd_trig = z.dcd_trig ? z.level > 0.1f : z.level > 4.0f;
}

// M17Demodulator.h:274-286 (update_dcd, dcd_on :243-257, dcd_off :259-265)
__device__ __forceinline__ void update_dcd(Scan& z, const ScanEnv& e)
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

// M17Demodulator.h:288-342
__device__ __forceinline__ void do_unlocked(Scan& z, const ScanEnv& e)
{
    int timing;
    if (z.missing < 1920) {
        z.missing += 1;
        int upd = do_sync_step(z, e, 0, kWordPreamble, 29.f, kLowest, &timing);
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
    int upd = do_sync_step(z, e, 1, kWordLsf, 31.f, -31.f, &timing);
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
    upd = do_sync_step(z, e, 2, kWordPacket, 31.f, -31.f, &timing);
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

// M17Demodulator.h:349-411
__device__ __forceinline__ void do_lsf_sync(Scan& z, const ScanEnv& e)
{
    if (z.prev_pos % 10 != z.si) return;
    float trig = sync_triggered(e.ring, z.prev_pos, z.limit, kWordPreamble, 29.f, kLowest);
    if ((double) trig > 0.1) {
        z.need_update = true;
        z.sync_count += 1;
        return;
    }
    trig = sync_triggered(e.ring, z.prev_pos, z.limit, kWordLsf, 31.f, -31.f);
    float bert = sync_triggered(e.ring, z.prev_pos, z.limit, kWordPacket, 31.f, -31.f);
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
__device__ __forceinline__ void do_stream_sync(Scan& z, const ScanEnv& e)
{
    if (sync_triggered(e.ring, z.prev_pos, z.limit, kWordEot, 31.f, kLowest) > 0.1f) {
        z.swt = SW_STREAM;
        z.state = ST_FRAME;
        z.eot = true;
        z.missing = 0;
        return;
    }
    int timing;
    int upd = do_sync_step(z, e, 1, kWordLsf, 31.f, -31.f, &timing);
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
__device__ __forceinline__ void do_packet_or_bert_sync(Scan& z, const ScanEnv& e, bool bert)
{
    int timing;
    int upd = do_sync_step(z, e, 2, kWordPacket, 31.f, -31.f, &timing);
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

// M17Demodulator.h:595-654 up to the decoder call (which decode_kernel performs)
__device__ __forceinline__ void do_frame(Scan& z, const ScanEnv& e, float filtered)
{
    int idx = z.prev_pos % 10;
    int diff = z.si - idx;
    if (diff == 5 || diff == -5) {
        // ClockRecovery.h:76-88
        // fmod(v, 10) is v itself for |v| < 10 and v -+ 10 (exact) for |v| < 20; libm otherwise
        const float v = z.ck_sest + z.ck_est * (float) z.ck_count;
        const float av = fabsf(v);
        double csw;
        if (av < 10.f) csw = (double) v;
        else if (av < 20.f) csw = (double) v - (v < 0 ? -10.0 : 10.0);
        else csw = fmod((double) v, 10.0);
        if (csw < 0.) csw += 10;
        else if (csw >= 10) csw -= 10;
        z.csi = wrap_index((int) (int8_t) round(csw));
        z.si = z.csi;
        return;
    }
    if (idx != z.si) return;

    float sample = filtered - z.offset;
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

__global__ void __launch_bounds__(kScanThreads)
scan_kernel(DevPtrs d, uint32_t head)
{
    extern __shared__ float smem[];
    float* ring_s = smem;                                  // [80][128]
    float* stage = smem + 80 * kScanThreads;               // [3][8][128]
    LlrTable* tab = reinterpret_cast<LlrTable*>(smem + (80 + 24) * kScanThreads);
    const int tid = threadIdx.x;
    if (tid < 43) {
        tab->thr[tid] = c_tab.llr_thr[tid];
        tab->val[tid] = (uint32_t) (uint8_t) c_tab.llr_val[tid][0] | ((uint32_t) (uint8_t) c_tab.llr_val[tid][1] << 8);
    }
    __syncthreads();
    const uint32_t s = blockIdx.x * kScanThreads + tid;
    if (s >= d.n_streams) return;

    ScanEnv e;
    e.llr_tab = tab;
    e.sb = d.st + st_index(s, 0);
    e.ring = ring_s + tid;
    e.sw_samples = d.sw_samples + s;
    e.llr = d.llr + s;
    e.np = d.np;
    uint32_t* sb = e.sb;

    uint32_t pos = sb[F_POS * 32];
    uint32_t fl = sb[F_FLAGS * 32];
    uint32_t avail = head - pos;
    if (avail == 0 || (fl & FL_PENDING)) {
        if (avail != 0) atomicAdd(&d.ctr->lagging, 1u);
        return;
    }
    uint32_t todo = avail < (uint32_t) kRoundCap ? avail : (uint32_t) kRoundCap;

    Scan z;
    z.re0 = __uint_as_float(sb[F_RE0 * 32]); z.im0 = __uint_as_float(sb[F_IM0 * 32]);
    z.re1 = __uint_as_float(sb[F_RE1 * 32]); z.im1 = __uint_as_float(sb[F_IM1 * 32]);
    z.l1 = __uint_as_float(sb[F_L1 * 32]); z.l2 = __uint_as_float(sb[F_L2 * 32]);
    z.level = __uint_as_float(sb[F_LEVEL * 32]);
    z.iir0 = __uint_as_float(sb[F_IIR0 * 32]); z.iir1 = __uint_as_float(sb[F_IIR1 * 32]);
    z.limit = __uint_as_float(sb[F_LIMIT * 32]);
    z.idev = __uint_as_float(sb[F_IDEV * 32]); z.offset = __uint_as_float(sb[F_OFFSET * 32]);
    z.evm = __uint_as_float(sb[F_EVM * 32]);
    z.ck_est = __uint_as_float(sb[F_CK_EST * 32]); z.ck_sest = __uint_as_float(sb[F_CK_SEST * 32]);
    z.ck_count = sb[F_CK_COUNT * 32];
    { uint32_t v = sb[F_CNT * 32]; z.count = v & 0xFFFF; z.init = v >> 16; }
    { uint32_t v = sb[F_IDX * 32]; z.corr_pos = v & 0xFF; z.framer_index = (v >> 8) & 0xFFFF; z.gap_k = v >> 24; }
    z.prev_pos = z.corr_pos == 0 ? 79 : z.corr_pos - 1;
    z.state = (fl >> FL_STATE_SHIFT) & 7; z.swt = (fl >> FL_SWT_SHIFT) & 3;
    z.si = (fl >> FL_SI_SHIFT) & 15; z.ssi = (fl >> FL_SSI_SHIFT) & 15; z.csi = (fl >> FL_CSI_SHIFT) & 15;
    z.dcd = fl & FL_DCD; z.dcd_trig = fl & FL_DCD_TRIG; z.need_reset = fl & FL_NEED_RESET;
    z.need_update = fl & FL_NEED_UPDATE; z.dev_reset = fl & FL_DEV_RESET; z.eot = fl & FL_EOT;
    z.pending = false;
    z.sync_count = (int) sb[F_SYNC_COUNT * 32]; z.missing = (int) sb[F_MISSING * 32];
    z.vcost = sb[F_VCOST * 32];
    z.sw = sb[F_SW * 32];

    for (int k = 0; k != 80; ++k) e.ring[k * kScanThreads] = d.corr_ring[(size_t) k * d.np + s];

    const float c0r = c_tab.sdft[0], c0i = c_tab.sdft[1], c1r = c_tab.sdft[2], c1i = c_tab.sdft[3];
    const uint32_t cmask = d.ring_chunks - 1;
    float* snap = d.fir_snap + s;

    while (todo != 0 && !z.pending) {
        // stage one 8-sample chunk of y, x and x 120 samples earlier (15 chunks back)
        const uint32_t c = pos >> 3;
        {
            const float4* py = reinterpret_cast<const float4*>(d.ring_y + ((size_t) (c & cmask) * d.np + s) * 8);
            const float4* px = reinterpret_cast<const float4*>(d.ring_x + ((size_t) (c & cmask) * d.np + s) * 8);
            const float4* po = reinterpret_cast<const float4*>(d.ring_x + ((size_t) ((c - 15) & cmask) * d.np + s) * 8);
            float4 a0 = py[0], a1 = py[1], b0 = px[0], b1 = px[1], o0 = po[0], o1 = po[1];
            float* st = stage + tid;
            st[0 * kScanThreads] = a0.x; st[1 * kScanThreads] = a0.y; st[2 * kScanThreads] = a0.z; st[3 * kScanThreads] = a0.w;
            st[4 * kScanThreads] = a1.x; st[5 * kScanThreads] = a1.y; st[6 * kScanThreads] = a1.z; st[7 * kScanThreads] = a1.w;
            st += 8 * kScanThreads;
            st[0 * kScanThreads] = b0.x; st[1 * kScanThreads] = b0.y; st[2 * kScanThreads] = b0.z; st[3 * kScanThreads] = b0.w;
            st[4 * kScanThreads] = b1.x; st[5 * kScanThreads] = b1.y; st[6 * kScanThreads] = b1.z; st[7 * kScanThreads] = b1.w;
            st += 8 * kScanThreads;
            st[0 * kScanThreads] = o0.x; st[1 * kScanThreads] = o0.y; st[2 * kScanThreads] = o0.z; st[3 * kScanThreads] = o0.w;
            st[4 * kScanThreads] = o1.x; st[5 * kScanThreads] = o1.y; st[6 * kScanThreads] = o1.z; st[7 * kScanThreads] = o1.w;
        }
        for (int j = pos & 7; j != 8 && todo != 0 && !z.pending; ++j, ++pos, --todo) {
            const float y_spec = stage[(0 + j) * kScanThreads + tid];
            const float x = stage[(8 + j) * kScanThreads + tid];
            const float xold = stage[(16 + j) * kScanThreads + tid];

            // ---- M17Demodulator::operator(), M17Demodulator.h:656-753 ----
            z.count++;
            {   // dcd(input): SlidingDFT.h:118-132, DataCarrierDetect.h:53-58
                float delta = x - xold;
                float a = z.re0 + delta, b = z.im0;
                z.re0 = a * c0r - b * c0i; z.im0 = a * c0i + b * c0r;
                a = z.re1 + delta; b = z.im1;
                z.re1 = a * c1r - b * c1i; z.im1 = a * c1i + b * c1r;
                z.l1 += z.re0 * z.re0 + z.im0 * z.im0;
                z.l2 += z.re1 * z.re1 + z.im1 * z.im1;
            }

            const bool initializing = z.init != 0;           // :667-673
            float filtered;
            if (initializing) {
                --z.init;
                filtered = y_spec;              // the filter has run on every sample since sample 0
            } else if (!z.dcd) {                // :675-689
                if (z.count % 384 == 0) {   // count_ in [1, 384] here
                    update_dcd(z, e);
                    dcd_update(z);
                    z.count = 0;
                }
                continue;
            } else if (z.gap_k < kTapsUsed) {
                filtered = slow_fir(snap, d.ring_x, d.np, cmask, s, pos, z.gap_k);
                z.gap_k++;
            } else {
                filtered = y_spec;
            }

            {   // correlator.sample(): Correlator.h:43-49, IirFilter.h:26-42
                float h2 = z.iir1, h1 = z.iir0;
                float h0 = fabsf(filtered);
                h0 -= -1.98148851f * h1;
                h0 -= 0.98165828f * h2;
                float r = 0.f;
                r += 4.24433681e-05f * h0;
                r += 8.48867363e-05f * h1;
                r += 4.24433681e-05f * h2;
                z.limit = r;
                z.iir0 = h0; z.iir1 = h1;
                e.ring[z.corr_pos * kScanThreads] = filtered;
                z.prev_pos = z.corr_pos;
                if (++z.corr_pos == 80) z.corr_pos = 0;
            }

            if (initializing) {
                z.count = 0;
                if (z.init == 0) {              // from the next sample on the filter is gated by !dcd_
                    z.gap_k = kTapsUsed;
                    take_snapshot(snap, d.ring_x, d.np, cmask, s, pos, kTapsUsed);
                }
                continue;
            }

            if (z.prev_pos % 10 == 0) {         // :695-709
                if (z.need_reset) {
                    Kalman k; kf_reset(k, (float) z.ssi);          // ClockRecovery.h:33-39
                    kf_store(sb, F_CK_X0, k);
                    z.ck_count = 0;
                    z.csi = z.ssi;
                    z.ck_est = 0.f;
                    z.need_reset = false;
                    z.si = z.ssi;
                } else if (z.need_update) {
                    float2 r = clock_update_index(sb, z.ssi, z.ck_count);
                    z.ck_sest = r.x;
                    z.csi = wrap_index((int) (int8_t) round((double) z.ck_sest));
                    z.ck_est = r.y;
                    z.ck_count = 0;
                    z.need_update = false;
                }
            }
            z.ck_count++;                       // clock_recovery(filtered_sample), :712

            switch (z.state) {                  // :714-740
            case ST_UNLOCKED: do_unlocked(z, e); break;
            case ST_LSF_SYNC: do_lsf_sync(z, e); break;
            case ST_STREAM_SYNC:
                z.sync_count += 1;
                if (z.sync_count >= 78) do_stream_sync(z, e);
                break;
            case ST_PACKET_SYNC:
                z.sync_count += 1;
                if (z.sync_count >= 78) do_packet_or_bert_sync(z, e, false);
                break;
            case ST_BERT_SYNC:
                z.sync_count += 1;
                if (z.sync_count >= 78) do_packet_or_bert_sync(z, e, true);
                break;
            case ST_SYNC_WAIT:                  // :582-593
                if (z.sync_count < 86) z.sync_count += 1;
                else { z.need_update = true; z.state = ST_FRAME; }
                break;
            default: do_frame(z, e, filtered); break;
            }

            if (z.count % 960 == 0) {           // :742-752
                const bool was = z.dcd;
                update_dcd(z, e);
                z.count = 0;
                dcd_update(z);
                if (was && !z.dcd) take_snapshot(snap, d.ring_x, d.np, cmask, s, pos, z.gap_k);
            }
        }
    }

    // write back
    sb[F_POS * 32] = pos;
    sb[F_RE0 * 32] = __float_as_uint(z.re0); sb[F_IM0 * 32] = __float_as_uint(z.im0);
    sb[F_RE1 * 32] = __float_as_uint(z.re1); sb[F_IM1 * 32] = __float_as_uint(z.im1);
    sb[F_L1 * 32] = __float_as_uint(z.l1); sb[F_L2 * 32] = __float_as_uint(z.l2);
    sb[F_LEVEL * 32] = __float_as_uint(z.level);
    sb[F_IIR0 * 32] = __float_as_uint(z.iir0); sb[F_IIR1 * 32] = __float_as_uint(z.iir1);
    sb[F_LIMIT * 32] = __float_as_uint(z.limit);
    sb[F_IDEV * 32] = __float_as_uint(z.idev); sb[F_OFFSET * 32] = __float_as_uint(z.offset);
    sb[F_EVM * 32] = __float_as_uint(z.evm);
    sb[F_CK_EST * 32] = __float_as_uint(z.ck_est); sb[F_CK_SEST * 32] = __float_as_uint(z.ck_sest);
    sb[F_CK_COUNT * 32] = z.ck_count;
    sb[F_CNT * 32] = (uint32_t) z.count | ((uint32_t) z.init << 16);
    sb[F_IDX * 32] = (uint32_t) z.corr_pos | ((uint32_t) z.framer_index << 8) | ((uint32_t) z.gap_k << 24);
    fl = ((uint32_t) z.state << FL_STATE_SHIFT) | ((uint32_t) z.swt << FL_SWT_SHIFT) | ((uint32_t) z.si << FL_SI_SHIFT) |
         ((uint32_t) z.ssi << FL_SSI_SHIFT) | ((uint32_t) z.csi << FL_CSI_SHIFT) |
         (z.dcd ? FL_DCD : 0) | (z.dcd_trig ? FL_DCD_TRIG : 0) | (z.need_reset ? FL_NEED_RESET : 0) |
         (z.need_update ? FL_NEED_UPDATE : 0) | (z.dev_reset ? FL_DEV_RESET : 0) | (z.eot ? FL_EOT : 0) |
         (z.pending ? FL_PENDING : 0);
    sb[F_FLAGS * 32] = fl;
    sb[F_SYNC_COUNT * 32] = (uint32_t) z.sync_count; sb[F_MISSING * 32] = (uint32_t) z.missing;
    sb[F_SW * 32] = z.sw;
    for (int k = 0; k != 80; ++k) d.corr_ring[(size_t) k * d.np + s] = e.ring[k * kScanThreads];
    if (pos != head) atomicAdd(&d.ctr->lagging, 1u);
}

void launch_scan(const DevPtrs& d, uint32_t head, cudaStream_t s)
{
    size_t smem = kScanSmem;
    scan_kernel<<<d.np / kScanThreads, kScanThreads, smem, s>>>(d, head);
}

// ---------------------------------------------------------------------------
// (C) frame decode
// ---------------------------------------------------------------------------

constexpr int kDecThreads = 128;

// CRC16.h:41-62 over the 30 LSF bytes (a column of dec_bytes); 0 means valid.
__device__ __noinline__ uint32_t crc16_lsf(const uint8_t* lsf, size_t np)
{
    uint32_t reg = c_tab.crc_init;
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

// One depunctured soft bit through the folded derandomise/deinterleave map.
__device__ __forceinline__ int soft_bit(const int16_t* llr, size_t np, int m, int stale)
{
    if (m == 0) return 0;
    if (m == kMapStale) return stale;
    int k = (m < 0 ? -m : m) - 1;
    int pair = llr[(size_t) (k >> 1) * np];
    int v = (k & 1) ? (int) (int8_t) (pair >> 8) : (int) (int8_t) (pair & 0xFF);
    return m < 0 ? -v : v;
}

// butterfly J: cost_[J][0] = +7 iff bit 2 of J, cost_[J][1] = +7 iff bit0 ^ bit1 of J
// (Viterbi.h:72-88 with polynomials 031, 027 and memory J << 1).
template <int J>
__device__ __forceinline__ void butterfly(const int (&m)[16], int (&n)[16], uint32_t& hist,
                                          int a0, int a1, int b0, int b1)
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

// Viterbi.h:181-236.  Writes out_bits(kind) bits MSB-first into out[] (to_byte_array,
// Util.h:300-318) and returns round(min_metric / 7.f).
__device__ __noinline__ uint32_t viterbi(int kind, const int16_t* llr, size_t np, int stale,
                                         uint16_t* hist_col /* stride kDecThreads */, uint8_t* out)
{
    const int16_t* map = c_tab.vmap[kind];
    const int n_in = map_len(kind), n_out = out_bits(kind), steps = n_in / 2;
    int m[16], n[16];
#pragma unroll
    for (int i = 0; i != 16; ++i) m[i] = 0x7fffffff / 2;
    m[0] = 0;
    for (int t = 0; t != steps; ++t) {
        int s0 = soft_bit(llr, np, map[2 * t], stale);
        int s1 = soft_bit(llr, np, map[2 * t + 1], stale);
        int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
        if (s0) { a0 = abs(7 - s0); a1 = abs(-7 - s0); }
        if (s1) { b0 = abs(7 - s1); b1 = abs(-7 - s1); }
        uint32_t h = 0;
        butterfly<0>(m, n, h, a0, a1, b0, b1); butterfly<1>(m, n, h, a0, a1, b0, b1);
        butterfly<2>(m, n, h, a0, a1, b0, b1); butterfly<3>(m, n, h, a0, a1, b0, b1);
        butterfly<4>(m, n, h, a0, a1, b0, b1); butterfly<5>(m, n, h, a0, a1, b0, b1);
        butterfly<6>(m, n, h, a0, a1, b0, b1); butterfly<7>(m, n, h, a0, a1, b0, b1);
        hist_col[t * kDecThreads] = (uint16_t) h;
#pragma unroll
        for (int i = 0; i != 16; ++i) m[i] = n[i];
    }
    int min_element = 0, min_cost = m[0];
#pragma unroll
    for (int i = 0; i != 16; ++i)
        if (m[i] < min_cost) { min_cost = m[i]; min_element = i; }
    uint32_t cost = (uint32_t) (long long) roundf((float) min_cost / 7.0f);

    const int n_bytes = (n_out + 7) / 8;
    for (int i = 0; i != n_bytes; ++i) out[i] = 0;
    int next = min_element, index = steps, o = n_out;
    for (int t = steps; t-- != 0 && o != 0;) {
        int v = (hist_col[t * kDecThreads] >> next) & 1;      // prevState_[next][v] = (next >> 1) | (v << 3)
        if (index-- <= n_out) { --o; out[o >> 3] |= (uint8_t) ((next & 1) << (7 - (o & 7))); }
        next = (next >> 1) | (v << 3);
    }
    return cost;
}

__device__ __noinline__ void emit_record(const DevPtrs& d, uint32_t s, uint32_t* sb, unsigned long long sample_pos,
                                         int type, int cost)
{
    uint32_t seq = sb[F_SEQ * 32];
    sb[F_SEQ * 32] = seq + 1;
    atomicAdd(&d.ctr->frames[type], 1ull);
    uint32_t slot = atomicAdd(&d.ctr->rec_count, 1u);
    if (slot >= d.rec_cap) { atomicAdd(&d.ctr->dropped, 1ull); return; }
    m17b_frame* r = d.recs + slot;
    r->stream = s; r->seq = seq; r->sample_pos = sample_pos; r->viterbi_cost = cost;
    r->type = (uint8_t) type; r->pad[0] = r->pad[1] = r->pad[2] = 0;
    const uint8_t* b = d.dec_bytes + s;
    for (int i = 0; i != 26; ++i) r->payload[i] = b[(size_t) i * d.np];
    for (int i = 0; i != 30; ++i) r->lsf[i] = b[(size_t) (26 + i) * d.np];
}

// unpack_lich, M17FrameDecoder.h:181-215, with Golay24::decode (Golay24.h:203-222).
// Returns false when a codeword is uncorrectable; lich[] holds whatever was
// assembled so far either way, like the reference's output_buffer.lich.
__device__ __noinline__ bool unpack_lich(const int16_t* llr, size_t np, uint8_t* lich)
{
    for (int i = 0; i != 6; ++i) lich[i] = 0;
    int index = 0;
    for (int i = 0; i != 4; ++i) {
        uint32_t codeword = 0;
        for (int j = 0; j != 24; ++j) {
            int v = soft_bit(llr, np, c_tab.lichmap[i * 24 + j], 0);
            codeword <<= 1;
            codeword |= (uint32_t) (v > 0);
        }
        uint32_t synd = (codeword >> 1) & 0xffffffu;          // syndrome(input >> 1), Golay24.h:102-112
        for (int b = 0; b != 12; ++b) { if (synd & 1) synd ^= 0xC75; synd >>= 1; }
        uint32_t decoded = codeword ^ (g_golay[synd] << 1);
        if (!(__popc(synd) < 3 || !(__popc(decoded) & 1))) return false;
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

__global__ void __launch_bounds__(kDecThreads)
decode_kernel(DevPtrs d, uint32_t head, unsigned long long head64)
{
    extern __shared__ uint16_t hist[];                     // [244][128] survivor decisions
    const int tid = threadIdx.x;
    const uint32_t s = blockIdx.x * kDecThreads + tid;
    if (s >= d.n_streams) return;
    uint32_t* sb = d.st + st_index(s, 0);
    uint32_t fl = sb[F_FLAGS * 32];
    if (!(fl & FL_PENDING)) return;

    const size_t np = d.np;
    const int16_t* llr = d.llr + s;
    uint8_t* bytes = d.dec_bytes + s;                      // rows 0..25 union payload, 26..55 lsf
    uint16_t* hcol = hist + tid;
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
        vcost = viterbi(0, llr, np, stale, hcol, out);
        stale = soft_bit(llr, np, c_tab.vmap[0][401], 0);              // depuncture_buffer.lsf[401]
        for (int i = 0; i != 30; ++i) bytes[(size_t) (26 + i) * np] = out[i];
        if (crc16_lsf(bytes + 26 * np, np) == 0) {
            // update_state, :113-136 (bit 111 = byte 13 bit 0, 109 = bit 2, 110 = bit 1)
            int b111 = out[13] & 1, b109 = (out[13] >> 2) & 1, b110 = (out[13] >> 1) & 1;
            if (b111) { if (b109) dec_state = DEC_STREAM; }
            else dec_state = ((b109 << 1) | b110) == 1 ? DEC_BASIC_PACKET : DEC_FULL_PACKET;
            emit_record(d, s, sb, sample_pos, M17B_FRAME_LSF, (int) vcost);
        } else {
            atomicAdd(&d.ctr->lsf_crc_fail, 1ull);
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
                atomicAdd(&d.ctr->golay_fail, 1ull);
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
                    } else {
                        vcost = 128;
                    }
                }
            }
        } else if (dec_state == DEC_STREAM) {
            // decode_stream, :276-289
            vcost = viterbi(1, llr, np, stale, hcol, out);
            for (int i = 0; i != 18; ++i) bytes[(size_t) i * np] = out[i];
            emit_record(d, s, sb, sample_pos, M17B_FRAME_STREAM, (int) vcost);
        } else {
            dec_state = DEC_LSF;
        }
    } else if (swt == SW_PACKET) {
        if (dec_state == DEC_BASIC_PACKET || dec_state == DEC_FULL_PACKET) {
            // decode_packet, :299-315
            vcost = viterbi(2, llr, np, stale, hcol, out);
            stale = soft_bit(llr, np, c_tab.vmap[2][401], 0);          // depuncture_buffer.packet[401]
            for (int i = 0; i != 26; ++i) bytes[(size_t) i * np] = out[i];
            emit_record(d, s, sb, sample_pos,
                        dec_state == DEC_BASIC_PACKET ? M17B_FRAME_BASIC_PACKET : M17B_FRAME_FULL_PACKET, (int) vcost);
            if (out[25] & 0x80) dec_state = DEC_LSF;
        } else {
            dec_state = DEC_LSF;
        }
    } else {
        // decode_bert, :264-274
        dec_state = DEC_BERT;
        vcost = viterbi(3, llr, np, stale, hcol, out);
        for (int i = 0; i != 25; ++i) bytes[(size_t) i * np] = out[i];
        emit_record(d, s, sb, sample_pos, M17B_FRAME_BERT, (int) vcost);
    }

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
    sb[F_FLAGS * 32] = fl;
}

void launch_decode(const DevPtrs& d, uint32_t head, unsigned long long head64, cudaStream_t s)
{
    size_t smem = sizeof(uint16_t) * 244 * kDecThreads;
    decode_kernel<<<d.np / kDecThreads, kDecThreads, smem, s>>>(d, head, head64);
}

// Per-device kernel attributes (dynamic shared memory above 48 KiB).
cudaError_t configure_kernels()
{
    cudaError_t e = cudaFuncSetAttribute(scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kScanSmem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int) (sizeof(uint16_t) * 244 * kDecThreads));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(fir_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int) (sizeof(float) * kFirStreams * kFirRow));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(fir_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int) (sizeof(float) * kFirStreams * kFirRow));
}

// Fresh-demodulator state for every stream: the member initialisers of
// M17Demodulator (M17Demodulator.h:148-178) and of its blocks; all other state is
// zero (cudaMemset by the host).
__global__ void init_state_kernel(DevPtrs d)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= d.np) return;
    uint32_t* sb = d.st + st_index(s, 0);
    for (int f = 0; f != NFIELDS; ++f) sb[f * 32] = 0;
    Kalman k; kf_reset(k, 0.f);                              // KalmanFilter.h:28-31
    kf_store(sb, F_CK_X0, k); kf_store(sb, F_DMIN_X0, k); kf_store(sb, F_DMAX_X0, k);
    sb[F_EVM * 32] = __float_as_uint(1.0f);                  // RunningStandardDeviation::S{1.0}
    sb[F_CNT * 32] = 1920u << 16;                            // `initializing`, M17Demodulator.h:659 (per stream)
    sb[F_IDX * 32] = (uint32_t) kTapsUsed << 24;             // no filter gap pending
    sb[F_FLAGS * 32] = FL_DEV_RESET | ((uint32_t) ST_UNLOCKED << FL_STATE_SHIFT) | ((uint32_t) SW_LSF << FL_SWT_SHIFT);
    sb[F_DEC * 32] = DEC_LSF;
}

void launch_init_state(const DevPtrs& d, cudaStream_t s)
{
    init_state_kernel<<<(d.np + 255) / 256, 256, 0, s>>>(d);
}

// ---------------------------------------------------------------------------
// FP32 pipe microbenchmark (roofline denominator for the matched filter)
// ---------------------------------------------------------------------------

template <int VARIANT>
__global__ void __launch_bounds__(256) fp32_peak_kernel(float* out, int iters, float seed)
{
    float a[16];
#pragma unroll
    for (int i = 0; i != 16; ++i) a[i] = seed + (float) (threadIdx.x + i);
    const float m = 0.999f, c = 1e-3f;
    for (int it = 0; it != iters; ++it) {
        if constexpr (VARIANT == 0) {
#pragma unroll
            for (int r = 0; r != 8; ++r)
#pragma unroll
                for (int i = 0; i != 16; ++i) a[i] = __fmaf_rn(a[i], 0.999f, 1e-3f);
        } else if constexpr (VARIANT == 1) {
#pragma unroll
            for (int r = 0; r != 8; ++r)
#pragma unroll
                for (int i = 0; i != 16; i += 2) {
                    unsigned long long v, mm, cc;
                    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[i]), "f"(a[i + 1]));
                    asm("mov.b64 %0, {%1, %1};" : "=l"(mm) : "f"(m));
                    asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
                    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(v) : "l"(v), "l"(mm), "l"(cc));
                    asm("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(a[i + 1]) : "l"(v));
                }
        } else if constexpr (VARIANT == 2) {
#pragma unroll
            for (int r = 0; r != 8; ++r)
#pragma unroll
                for (int i = 0; i != 16; ++i) a[i] = __fadd_rn(__fmul_rn(a[i], 0.999f), 1e-3f);
        } else {
            // register-operand FFMA (no immediates): multiplier and addend live in registers
            float mr = out[0] * 0.f + m, cr = out[1] * 0.f + c;
#pragma unroll
            for (int r = 0; r != 8; ++r)
#pragma unroll
                for (int i = 0; i != 16; ++i) a[i] = __fmaf_rn(a[i], mr, cr);
        }
    }
    float sum = 0;
#pragma unroll
    for (int i = 0; i != 16; ++i) sum += a[i];
    if (sum == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
}

double run_fp32_peak(int variant, int sm_count, cudaStream_t s)
{
    float* buf = nullptr;
    if (cudaMalloc(&buf, 1 << 20) != cudaSuccess) return -1.0;
    cudaMemsetAsync(buf, 0, 1 << 20, s);
    const int iters = 4096, blocks = sm_count * 8, threads = 256;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep != 5; ++rep) {
        cudaEventRecord(e0, s);
        switch (variant) {
        case 0: fp32_peak_kernel<0><<<blocks, threads, 0, s>>>(buf, iters, 1.0f); break;
        case 1: fp32_peak_kernel<1><<<blocks, threads, 0, s>>>(buf, iters, 1.0f); break;
        case 2: fp32_peak_kernel<2><<<blocks, threads, 0, s>>>(buf, iters, 1.0f); break;
        default: fp32_peak_kernel<3><<<blocks, threads, 0, s>>>(buf, iters, 1.0f); break;
        }
        cudaEventRecord(e1, s);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        // flop counted as 2 per multiply-add pair in every variant
        double flop = 2.0 * 16 * 8 * (double) iters * threads * (double) blocks;
        double tf = flop / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(buf);
    return best;
}

} // namespace m17b
