// m17b_kernels.cu -- hand-written sm_100a kernels of the M17 demodulation path.
//
//   fir_kernel     (A) int16 -> float scaling (apps/m17-demod.cpp:489) + 150-tap RRC
//                      matched filter (FirFilter.h:28-43, taps M17Demodulator.h:79-118)
//                      for every stream and every new sample; FP32 pipe.
//   dcd_kernel     (A2) the carrier detect's two-bin sliding DFT (SlidingDFT.h:118-132),
//                      all streams in lock-step over a submit's new samples.
//   scan_kernel    (B) per-stream serial part of M17Demodulator::operator()
//                      (M17Demodulator.h:656-753): limit biquad, carrier-detect cadence,
//                      sync windows, Kalman clock recovery, deviation / offset estimator,
//                      4-FSK soft slicer, framer.  One thread per stream, every lane at its
//                      own stream position.
//   fixup_kernel   (B2) what a stream asks for when it cannot go on alone: the matched
//                      filter outputs after a carrier-detect gap (stale history), and the
//                      sync-word search of ST_UNLOCKED (Correlator.h:43-64,146-157) over the
//                      samples ahead of it, one thread per sample.
//   decode_kernel  (C) M17FrameDecoder::operator() (M17FrameDecoder.h:353-392):
//                      derandomise + deinterleave + depuncture folded into one index
//                      map, soft Viterbi, Golay24 LICH, CRC16, decoder state machine,
//                      frame-record emission.
//
// Compiled with -fmad=false: every float operation below is a separately rounded
// IEEE operation unless it is an explicit __fmaf_rn.
#include "m17b_dev.cuh"

#include <cstdio>
#include <cstdlib>
#include <algorithm>

namespace m17b {

// ---------------------------------------------------------------------------
// constants
// ---------------------------------------------------------------------------

__constant__ ConstTables c_tab;
__device__ uint32_t g_golay[2048];

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
// (tests/test_gpu_parity.py::test_scaling_all_int16_values_bit_exact).
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
constexpr int kFirOut = 16;         // outputs per thread and pass
constexpr int kFirPass = kFirWarps * kFirOut;      // 128 positions per pass
constexpr int kFirPasses = 2;
constexpr int kFirTile = kFirPass * kFirPasses;    // 256 positions per CTA
constexpr int kFirHalo = 160;
constexpr int kFirCols = kFirTile + kFirHalo;      // 416 staged samples per stream
constexpr int kFirRow = kFirCols + 1;              // 417: odd stride -> conflict-free columns
constexpr int kFirLoop = 9;                        // taps 0..143 in a rolled loop of 16, 144..148 unrolled

// Two accumulators in one 64-bit register pair.  Exact mode adds two separately rounded products
// with one packed FADD2 (add.rn.f32x2): half the issue slots of two FADDs, same rounding per half.
// (A packed mul.rn.f32x2 feeding it is NOT used: ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2
// into FFMA2 even under --fmad=false, which would change the rounding.)
struct Acc2 { unsigned long long v; };

__device__ __forceinline__ void acc2_add(Acc2& a, float p0, float p1)
{
    unsigned long long p;
    asm("mov.b64 %0, {%1, %2};" : "=l"(p) : "f"(p0), "f"(p1));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(a.v) : "l"(a.v), "l"(p));
}

__device__ __forceinline__ void acc2_fma(Acc2& a, float x0, float x1, float tap)
{
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v));
    lo = __fmaf_rn(x0, tap, lo);
    hi = __fmaf_rn(x1, tap, hi);
    asm("mov.b64 %0, {%1, %2};" : "=l"(a.v) : "f"(lo), "f"(hi));
}

// One tap for 16 consecutive outputs held in acc[] (pairs); w[] is a 16-entry register window: slot
// (k & 15) holds sample t0 + k.  Tap index I = 16 * it + K; rowp = row - 16 * it.
template <int K, int KEND, bool FMA>
struct FirTaps {
    __device__ __forceinline__ static void run(Acc2 (&acc)[kFirOut / 2], float (&w)[kFirOut], const float* rowp, const float* tp)
    {
        w[(16 - K) & 15] = rowp[-K];               // sample t0 - I replaces t0 + 16 - I
        const float tap = tp[K];
#pragma unroll
        for (int r = 0; r != kFirOut; r += 2) {
            float x0 = w[(r - K) & 15], x1 = w[(r + 1 - K) & 15];
            if constexpr (FMA) acc2_fma(acc[r / 2], x0, x1, tap);
            else acc2_add(acc[r / 2], __fmul_rn(x0, tap), __fmul_rn(x1, tap));
        }
        if constexpr (K + 1 < KEND) FirTaps<K + 1, KEND, FMA>::run(acc, w, rowp, tp);
    }
};

template <bool FMA>
__global__ void __launch_bounds__(kFirWarps * 32)
fir_kernel(DevPtrs d, const int16_t* __restrict__ in, size_t stride, uint32_t n_new,
           const int16_t* __restrict__ tail, uint32_t head, int invert, int aligned)
{
    extern __shared__ float xs[];                  // [32][kFirRow]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t s0 = blockIdx.x * kFirStreams;
    const uint32_t b0 = head & ~7u;
    const uint32_t tb = b0 + blockIdx.y * kFirTile;          // first position of this tile
    const int rel0 = (int) (tb - head) - kFirHalo;            // input index of tile column 0

    // stage [tb-160, tb+256) of 32 streams, scaled to float: lane = stream, 16 samples (one
    // 32-byte sector of int16) per thread and step
    {
        const uint32_t s = s0 + lane;
        const bool live = s < d.n_streams;
        float* dst = xs + lane * kFirRow;
        const int16_t* src_in = in + (size_t) s * stride;
        const int16_t* src_tail = tail + (size_t) s * kTailLen + kTailLen;
        for (int g = warp; g < kFirCols / 16; g += kFirWarps) {
            const int t = g * 16, rel = rel0 + t;
            int v[16];
            if (aligned && live && (rel + 16 <= 0 ? rel >= -kTailLen : (rel >= 0 && (uint32_t) rel + 16u <= n_new))) {
                const uint4* p = reinterpret_cast<const uint4*>(rel < 0 ? src_tail + rel : src_in + rel);
                uint4 q0 = p[0], q1 = p[1];
                const uint32_t u[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
                for (int k = 0; k != 8; ++k) { v[2 * k] = (int) (int16_t) (u[k] & 0xFFFF); v[2 * k + 1] = (int) (int16_t) (u[k] >> 16); }
            } else {
#pragma unroll
                for (int k = 0; k != 16; ++k) {
                    int r = rel + k, x = 0;
                    if (live) {
                        if (r >= 0) { if ((uint32_t) r < n_new) x = src_in[r]; }
                        else if (r >= -kTailLen) x = src_tail[r];
                    }
                    v[k] = x;
                }
            }
#pragma unroll
            for (int k = 0; k != 16; ++k) {
                int x = v[k];
                if (invert) x = (int) (int16_t) (-x);                 // int16 `sample *= -1`
                dst[t + k] = scale_sample(x);
            }
        }
    }
    __syncthreads();

    const uint32_t s = s0 + lane;
    const uint32_t lmask = d.ring_lines - 1;
#pragma unroll 1
    for (int pass = 0; pass != kFirPasses; ++pass) {
        const int o0 = pass * kFirPass + warp * kFirOut;              // first output of this thread in the tile
        if ((int32_t) (tb + (uint32_t) o0 - (head + n_new)) >= 0) break;   // nothing new from here on (warp-uniform; positions wrap)
        const float* row = xs + lane * kFirRow + kFirHalo + o0;        // sample t0
        Acc2 acc2[kFirOut / 2];
        float acc[kFirOut], w[kFirOut];
#pragma unroll
        for (int r = 0; r != kFirOut; ++r) w[r] = row[r];
#pragma unroll
        for (int r = 0; r != kFirOut / 2; ++r) acc2[r].v = 0ull;
#pragma unroll 1
        for (int it = 0; it != kFirLoop; ++it)
            FirTaps<0, 16, FMA>::run(acc2, w, row - 16 * it, c_tab.taps + 16 * it);
        FirTaps<0, kTapsUsed - 16 * kFirLoop, FMA>::run(acc2, w, row - 16 * kFirLoop, c_tab.taps + 16 * kFirLoop);
#pragma unroll
        for (int r = 0; r != kFirOut / 2; ++r)
            asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[2 * r]), "=f"(acc[2 * r + 1]) : "l"(acc2[r].v));

        // 16 outputs = two 8-sample blocks; a block never straddles a 32-sample ring line.  Besides y: the raw
        // sample as int16 (exact inverse of the scaling), for the rare paths that rebuild filter history.
#pragma unroll
        for (int k = 0; k != kFirOut / 8; ++k) {
            const uint32_t p = tb + (uint32_t) o0 + 8u * k;
            const size_t o = ((size_t) ((p >> 5) & lmask) * d.np + s) * kLine + (p & (kLine - 1));
            float4* py = reinterpret_cast<float4*>(d.ring_y + o);
            uint32_t raw[4];
#pragma unroll
            for (int i = 0; i != 4; ++i) {
                int lo = __float2int_rn(__fmul_rn(row[8 * k + 2 * i], 41067.0f));
                int hi = __float2int_rn(__fmul_rn(row[8 * k + 2 * i + 1], 41067.0f));
                raw[i] = ((uint32_t) lo & 0xFFFFu) | ((uint32_t) hi << 16);
            }
            if ((int) (p - head) < 0) {
                // the block that contains `head` of an unaligned submit: its first samples belong to the
                // previous submit and may since have been corrected by fixup_kernel for a stream that has
                // not consumed them yet (M17B_STREAMING) -- write the new samples only
#pragma unroll
                for (int i = 0; i != 8; ++i)
                    if ((int) (p + (uint32_t) i - head) >= 0) {
                        d.ring_y[o + i] = acc[8 * k + i];
                        d.ring_x[o + i] = (int16_t) ((i & 1) ? (raw[i >> 1] >> 16) : (raw[i >> 1] & 0xFFFFu));
                    }
                continue;
            }
            py[0] = make_float4(acc[8 * k + 0], acc[8 * k + 1], acc[8 * k + 2], acc[8 * k + 3]);
            py[1] = make_float4(acc[8 * k + 4], acc[8 * k + 5], acc[8 * k + 6], acc[8 * k + 7]);
            *reinterpret_cast<uint4*>(d.ring_x + o) = make_uint4(raw[0], raw[1], raw[2], raw[3]);
        }
    }
}

void launch_fir(const DevPtrs& d, const int16_t* in, size_t stride, uint32_t n_new, const int16_t* tail,
                uint32_t head, bool fma, bool invert, cudaStream_t s)
{
    uint32_t b0 = head & ~7u;
    uint32_t span = head + n_new - b0;
    dim3 grid(d.np / kFirStreams, (span + kFirTile - 1) / kFirTile);
    size_t smem = sizeof(float) * kFirStreams * kFirRow;
    // 16-byte loads need the chunk to start on a ring-chunk boundary and 16-byte aligned rows
    int aligned = (head % 8 == 0) && (stride % 8 == 0) && (reinterpret_cast<uintptr_t>(in) % 16 == 0) &&
                  (reinterpret_cast<uintptr_t>(tail) % 16 == 0);
    if (fma) fir_kernel<true><<<grid, kFirWarps * 32, smem, s>>>(d, in, stride, n_new, tail, head, invert, aligned);
    else fir_kernel<false><<<grid, kFirWarps * 32, smem, s>>>(d, in, stride, n_new, tail, head, invert, aligned);
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
    size_t o = ((size_t) ((i >> 5) & (d.ring_lines - 1)) * d.np + s) * kLine + (i & (kLine - 1));
    y[(size_t) s * n + i] = d.ring_y[o];
    if (x) x[(size_t) s * n + i] = scale_sample(d.ring_x[o]);
}

void launch_gather(const DevPtrs& d, uint32_t n, float* y, float* x, cudaStream_t s)
{
    dim3 grid((n + 255) / 256, d.n_streams);
    gather_ring_kernel<<<grid, 256, 0, s>>>(d, n, y, x);
}

// ---------------------------------------------------------------------------
// (B) per-stream scan and (C) frame decode: the per-lane logic lives in m17b_lane.cuh
// ---------------------------------------------------------------------------

} // namespace m17b

#include "m17b_lane.cuh"

namespace m17b {

constexpr int kScanThreads = 128;

// dcd_lane's sample source on the device: x[n] - x[n-120] from the submit's int16 input and the raw history tail.
// The two lanes of a stream (one per DFT bin) share the work of a whole block inside the submit: each loads,
// scales and subtracts four of its eight samples (8-byte loads; x[n-120] is 15 blocks back, in the same row or in
// the tail, never straddling the two) and hands them to its partner by shuffle.  Blocks that stick out of the
// submit, and every block of a submit whose rows are not 16-byte aligned, take the scalar path.
struct DcdSrc {
    const int16_t* row;        // the stream's new samples: row[0 .. n_new)
    const int16_t* trow;       // one past the stream's raw history: trow[-160 .. 0)
    uint32_t head_old, n_new, mask;
    int half, invert, aligned;
    int rel0;
    bool fast;
    uint2 cur, old;
    __device__ __forceinline__ float sx(int v) const { return scale_i16(invert ? (int) (int16_t) (-v) : v); }
    __device__ __forceinline__ void request(uint32_t base)
    {
        rel0 = (int) (base - head_old);
        fast = aligned && rel0 >= 0 && (uint32_t) rel0 + 8u <= n_new;
        if (fast) {
            const int r = rel0 + 4 * half, ro = r - kDftN;
            cur = *reinterpret_cast<const uint2*>(row + r);
            old = *reinterpret_cast<const uint2*>((ro >= 0 ? row : trow) + ro);
        }
    }
    __device__ __forceinline__ void take(float (&dl)[8])
    {
        if (fast) {
            const uint32_t c[2] = {cur.x, cur.y}, o[2] = {old.x, old.y};
            float v[4], w[4];
#pragma unroll
            for (int i = 0; i != 4; ++i) {
                const int xc = (int) (int16_t) ((i & 1) ? (c[i >> 1] >> 16) : (c[i >> 1] & 0xFFFFu));
                const int xo = (int) (int16_t) ((i & 1) ? (o[i >> 1] >> 16) : (o[i >> 1] & 0xFFFFu));
                v[i] = __fsub_rn(sx(xc), sx(xo));
            }
#pragma unroll
            for (int i = 0; i != 4; ++i) w[i] = __shfl_xor_sync(mask, v[i], 1);
#pragma unroll
            for (int i = 0; i != 4; ++i) { dl[i] = half ? w[i] : v[i]; dl[4 + i] = half ? v[i] : w[i]; }
        } else {
#pragma unroll
            for (int j = 0; j != 8; ++j) {
                const int r = rel0 + j, ro = r - kDftN;
                float v = 0.f;
                if (r >= 0 && (uint32_t) r < n_new) v = __fsub_rn(sx(row[r]), sx(ro >= 0 ? row[ro] : trow[ro]));
                dl[j] = v;
            }
        }
    }
};

// Carrier-detect pre-pass (dcd_lane) over the new samples of a submit, all lanes at the same position, two lanes (one
// per DFT bin) per stream.  Bound: FP32 issue; reads the int16 input once from DRAM (2 B/sample).
__global__ void __launch_bounds__(128)
dcd_kernel(DevPtrs d, const int16_t* __restrict__ in, size_t stride, const int16_t* __restrict__ tail, uint32_t head_old, uint32_t n_new,
           unsigned long long head64_old, int invert, int aligned)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;      // lane pair = the two bins of one stream
    if ((t >> 1) >= d.n_streams) return;
    const uint32_t s = t >> 1;
    DcdSrc src;
    src.row = in + (size_t) s * stride;
    src.trow = tail + (size_t) s * kTailLen + kTailLen;
    src.head_old = head_old; src.n_new = n_new; src.mask = __activemask();
    src.half = (int) (t & 1u); src.invert = invert; src.aligned = aligned;
    dcd_lane(d, s, (int) (t & 1u), head_old, n_new, head64_old, src);
}

void launch_dcd(const DevPtrs& d, const int16_t* in, size_t stride, const int16_t* tail, uint32_t head_old, uint32_t n_new,
                unsigned long long head64_old, bool invert, cudaStream_t s)
{
    // 8-byte loads need the submit to start on a ring-block boundary and 16-byte aligned rows (like the filter's)
    const int aligned = (head_old % 8 == 0) && (stride % 8 == 0) && (reinterpret_cast<uintptr_t>(in) % 16 == 0) &&
                        (reinterpret_cast<uintptr_t>(tail) % 16 == 0);
    dcd_kernel<<<(2 * d.n_streams + 127) / 128, 128, 0, s>>>(d, in, stride, tail, head_old, n_new, head64_old, invert ? 1 : 0, aligned);
}

// Which streams have samples to walk this round, and how: the lanes of a warp run in lock-step, so a warp moves at
// the pace of its slowest lane, and a stream that is acquiring (start-up, sync-word search, preamble: the state
// machine acts on every symbol or every sample) is an order of magnitude slower per sample than one that is
// receiving frames (one slicer decision per symbol).  With streams mapped to lanes by index nearly every warp
// carries one of the few slow ones.  classify_kernel therefore sorts the streams that have work into kClasses lists
// and scan_kernel's warps each take 32 entries of ONE list (per-stream state is then gathered rather than
// coalesced: a few hundred bytes per stream and round).
__global__ void __launch_bounds__(256)
classify_kernel(DevPtrs d, uint32_t head, uint32_t class_mask /* bit c: list class c this round */, int defer)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    int cls = -1;
    bool lag = false, dfr = false;
    if (s < d.n_streams) {
        const volatile uint32_t* sb = d.st + st_index(s, 0);
        // the flags word first: a stream whose flags make it this round's is complete (scan_lane / decode_lane
        // publish the flags last); one that another chain still works on shows flags that are not
        const uint32_t fl = sb[F_FLAGS * 32];
        __threadfence();
        const uint32_t pos = sb[F_POS * 32];
        if ((int32_t) (head - pos) > 0) {                       // (a stream may be past this chain's head: the other chain's rounds may belong to a later submit)
            if (fl & (FL_PENDING | FL_FIXUP)) lag = true;       // waits for decode_kernel / fixup_kernel: not at the head yet
            else {
                const int st = (int) ((fl >> FL_STATE_SHIFT) & 7);
                const bool receiving = (fl & FL_DCD) && (sb[F_CNT * 32] >> 16) == 0 && (sb[F_IDX * 32] >> 24) >= (uint32_t) kTapsUsed &&
                                       (st == ST_FRAME || st == ST_STREAM_SYNC || st == ST_PACKET_SYNC || st == ST_BERT_SYNC || st == ST_SYNC_WAIT);
                cls = receiving ? 0 : 1;
                if (!((class_mask >> cls) & 1u)) cls = -1;      // the other chain's
                else if (defer && defers_at_once(sb, fl, pos, head)) { cls = -1; dfr = true; }
            }
        }
    }
    const uint32_t lane = threadIdx.x & 31u;
#pragma unroll
    for (int c = 0; c != kClasses; ++c) {
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, cls == c);
        if (m) {
            uint32_t base = 0;
            if (lane == (uint32_t) (__ffs((int) m) - 1)) base = atomicAdd(&d.ctr->cls[c].scan_count, (unsigned) __popc(m));
            base = __shfl_sync(0xFFFFFFFFu, base, __ffs((int) m) - 1);
            if (cls == c) d.scan_list[(size_t) c * d.np + base + (uint32_t) __popc(m & ((1u << lane) - 1u))] = s;
        }
    }
    const uint32_t md = __ballot_sync(0xFFFFFFFFu, dfr);
    if (md && lane == 0) atomicAdd(&d.ctr->cls[(class_mask & 1u) ? 0 : 1].deferred, (unsigned) __popc(md));
    const uint32_t ml = __ballot_sync(0xFFFFFFFFu, lag);
    if (ml && lane == 0) atomicAdd(&d.ctr->cls[(class_mask & 1u) ? 0 : 1].lagging, (unsigned) __popc(ml));
}

void launch_classify(const DevPtrs& d, uint32_t head, uint32_t class_mask, bool defer, cudaStream_t s)
{
    classify_kernel<<<(d.np + 255) / 256, 256, 0, s>>>(d, head, class_mask, defer ? 1 : 0);
}

// The acquiring streams (class 1) are few and each is in a different part of the state machine: 32 of them in a
// warp run one after the other (2-3 of 32 lanes active per instruction, measured).  They are spread over kSpread
// times as many warps instead, every kSpread-th lane active, which shortens a warp's serial chain.  What bounds a
// round of theirs is the slowest single lane (a whole preamble in ST_LSF_SYNC: correlations and estimator updates on
// every symbol), so the gain is modest: configs[3] 23.0 -> 21.6 ms per step with 4, 22.1 with 8; the steady
// configs[2] run is unchanged (12.2-12.4).
// Only while the spread-out lanes still fit the machine in one wave (148 SMs x 16 resident warps of this kernel): with
// more acquiring streams than that (configs[3] at 131,072 channels: 40,000 of them) the extra warps are a second wave and
// cost more than the shorter chains save (39.2 against 35.5 ms per step), so the factor falls to 2 and to 1.
constexpr int kSpread = 4;
constexpr uint32_t kSpreadThreads = 148u * 16u * 32u;
__device__ __forceinline__ uint32_t spread_for(uint32_t n)
{
    return n * 4u <= kSpreadThreads ? 4u : n * 2u <= kSpreadThreads ? 2u : 1u;
}

__global__ void __launch_bounds__(kScanThreads)
scan_kernel(DevPtrs d, uint32_t head, unsigned long long head64, int defer, uint32_t guard, int cls, int split)
{
    __shared__ LlrTable tab;                               // lanes look up different entries: shared, not constant memory
    const int tid = threadIdx.x;
    const uint32_t t = blockIdx.x * kScanThreads + tid;
    const uint32_t n = d.ctr->cls[cls].scan_count;
    const uint32_t spread = cls ? spread_for(n) : 1u;
    const uint32_t g = t / spread;                         // thread takes the g-th entry of the class's list
    if (blockIdx.x * kScanThreads / spread >= n) return;   // CTA-uniform
    for (int i = tid; i < 48; i += kScanThreads) tab.thr[i] = llr_thr_padded(c_tab.llr_thr, i);
    for (int i = tid; i < 43; i += kScanThreads) {
        tab.val[i] = (uint32_t) (uint8_t) c_tab.llr_val[i][0] | ((uint32_t) (uint8_t) c_tab.llr_val[i][1] << 8);
    }
    __syncthreads();
    if (g >= n || t % spread != 0) return;
    const uint32_t s = d.scan_list[(size_t) cls * d.np + g];
    scan_lane(d, make_io(d, s, &tab, cls, split != 0), head, head64, true, defer != 0, guard);
}

void launch_scan(const DevPtrs& d, uint32_t head, unsigned long long head64, bool defer, uint32_t guard, int cls, bool split, cudaStream_t s)
{
    scan_kernel<<<d.np / kScanThreads * (cls ? kSpread : 1), kScanThreads, 0, s>>>(d, head, head64, defer ? 1 : 0, guard, cls, split ? 1 : 0);
}

// Thread g takes the g-th entry of the per-kind lists the scan kernel filled, each list padded to
// whole warps: all lanes of a warp decode the same kind of frame (same Viterbi length and map).
__global__ void __launch_bounds__(kDecThreads)
decode_kernel(DevPtrs d, uint32_t head, unsigned long long head64, int cls)
{
    extern __shared__ uint16_t hist[];                     // [244][128] survivor decisions
    uint32_t g = blockIdx.x * kDecThreads + threadIdx.x;
    uint32_t s = 0xFFFFFFFFu;
#pragma unroll
    for (int k = 0; k != 5; ++k) {
        const uint32_t n = d.ctr->cls[cls].dec_count[k], padded = (n + 31u) & ~31u;
        if (s == 0xFFFFFFFFu) {
            if (g < padded) { s = g < n ? d.dec_list[(size_t) (cls * 5 + k) * d.np + g] : 0xFFFFFFFEu; }
            else g -= padded;
        }
    }
    if (s >= d.n_streams) return;
    decode_lane(d, s, hist + threadIdx.x, head, head64);
}

// What a stream asks for when it cannot go on alone (FL_FIXUP), three launches per round:
//
// fixup_kernel (one CTA per request): matched-filter fix-up after a carrier-detect gap
//     (M17Demodulator.h:675-689: the filter is not clocked while !dcd_, so its first 149 outputs after it
//     resumes see stale history; fir_kernel computed them for contiguous history).  Thread k recomputes
//     output k in the reference's tap order from the snapshot and the scaled inputs staged in shared memory.
//
// The sync-word search of a stream in ST_UNLOCKED with the carrier detect on (search_lane in m17b_lane.cuh is
// the serial statement of the same pass): do_unlocked() evaluates up to two correlations and their thresholds
// on EVERY sample (M17Demodulator.h:288-342).  That is feed-forward once limit_ is known, so it is split into
// search_limit_kernel (one thread per request): the limit biquad over [F_POS, head), the only recurrence
//     (~12 dependent cycles per sample), leaving its history at every block end (ckpt_ring), and
// search_mask_kernel (one thread per request and 8-sample block): limit_ for the block's samples from the
//     block before's history, three 8-tap correlations per sample straight from the sample ring and
//     SyncWord::triggered's comparison -> one mask word per block.
// The stream then jumps from trigger to trigger through the masks (scan_lane's search walk) and adopts the
// limit-filter history where it leaves the search.
constexpr int kFixThreads = 160;

__global__ void __launch_bounds__(kFixThreads)
fixup_kernel(DevPtrs d, uint32_t head, int cls)
{
    __shared__ float xs[kTapsUsed], snap[kTapsUsed];
    const uint32_t n_req = d.ctr->cls[cls].fix_count;
    const uint32_t* const fix_list = d.fix_list + (size_t) cls * d.np;
    const uint32_t cmask = d.ring_lines - 1;
    const int t = threadIdx.x;
    for (uint32_t q = blockIdx.x; q < n_req; q += gridDim.x) {
        const uint32_t s = fix_list[q];
        uint32_t* sb = d.st + st_index(s, 0);
        const uint32_t R = sb[F_GAPR * 32];
        const int done = (int) sb[F_GAPDONE * 32];
        const uint32_t avail = head - R;
        const int upto = avail < (uint32_t) kTapsUsed ? (int) avail : kTapsUsed;
        if (done >= upto) continue;                         // CTA-uniform
        if (t < kTapsUsed) {
            snap[t] = d.fir_snap[(size_t) t * d.np + s];
            xs[t] = t < upto ? scale_i16(d.ring_x[ring_off(R + (uint32_t) t, cmask, d.np, s)]) : 0.f;
        }
        __syncthreads();
        if (t >= done && t < upto) {
            float result = 0.0f;
            for (int i = 0; i != kTapsUsed; ++i) {
                float v = i <= t ? xs[t - i] : snap[i - t - 1];
                result = __fadd_rn(result, __fmul_rn(v, c_tab.taps[i]));
            }
            d.ring_y[ring_off(R + (uint32_t) t, cmask, d.np, s)] = result;
        }
        __syncthreads();
        if (t == 0) sb[F_GAPDONE * 32] = (uint32_t) upto;
    }
}

__global__ void __launch_bounds__(128)
search_limit_kernel(DevPtrs d, uint32_t head, int cls)
{
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= d.ctr->cls[cls].fix_count) return;
    const uint32_t s = d.fix_list[(size_t) cls * d.np + q];
    const uint32_t* sb = d.st + st_index(s, 0);
    if (!wants_search(sb, head)) return;
    const uint32_t lmask = d.ring_lines - 1;
    const uint32_t from = sb[F_POS * 32];
    head = search_end(from, head);
    Biquad b;
    b.h1 = u2f(sb[F_IIR0 * 32]); b.h2 = u2f(sb[F_IIR1 * 32]);
    // whole ring lines (32 samples = four blocks) from the one that contains `from`; the next line is requested while
    // this one's 32 filter steps run, the one after that is pulled into L2 (the recurrence is ~12 dependent cycles per
    // sample, a line's worth of it about as long as an L2 round trip)
    uint32_t base = from & ~31u;
    float4 cur[8], nxt[8];
    {
        const float4* py = reinterpret_cast<const float4*>(d.ring_y + ring_off(base, lmask, d.np, s));
#pragma unroll
        for (int i = 0; i != 8; ++i) cur[i] = py[i];
    }
    for (; (int32_t) (head - base) > 0; base += 32u) {
        {
            const float4* py = reinterpret_cast<const float4*>(d.ring_y + ring_off(base + 32u, lmask, d.np, s));
#pragma unroll
            for (int i = 0; i != 8; ++i) nxt[i] = py[i];
            asm volatile("prefetch.global.L2 [%0];" :: "l"(d.ring_y + ring_off(base + 96u, lmask, d.np, s)));
        }
#pragma unroll
        for (int k = 0; k != 4; ++k) {
            const uint32_t b8 = base + 8u * (uint32_t) k;
            if ((int32_t) (b8 + 8u - from) <= 0 || (int32_t) (head - b8) <= 0) continue;     // before the span / past its end
            const float y[8] = {cur[2 * k].x, cur[2 * k].y, cur[2 * k].z, cur[2 * k].w, cur[2 * k + 1].x, cur[2 * k + 1].y, cur[2 * k + 1].z, cur[2 * k + 1].w};
            const uint32_t left = head - b8;
            if ((int32_t) (b8 - from) >= 0 && left >= 8u) {
#pragma unroll
                for (int j = 0; j != 8; ++j) limit_step(b, y[j]);
            } else {
#pragma unroll
                for (int j = 0; j != 8; ++j)
                    if ((int32_t) (b8 + (uint32_t) j - from) >= 0 && (uint32_t) j < left) limit_step(b, y[j]);
            }
            if (left >= 8u) d.ckpt_ring[blk_off(b8, lmask, d.np, s)] = make_float2(b.h1, b.h2);
        }
#pragma unroll
        for (int i = 0; i != 8; ++i) cur[i] = nxt[i];
    }
}

constexpr int kMaskThreads = 128;

__global__ void __launch_bounds__(kMaskThreads)
search_mask_kernel(DevPtrs d, uint32_t head_all, int cls)
{
    const uint32_t n_req = d.ctr->cls[cls].fix_count;
    const uint32_t* const fix_list = d.fix_list + (size_t) cls * d.np;
    const uint32_t lmask = d.ring_lines - 1;
    for (uint32_t q = blockIdx.y; q < n_req; q += gridDim.y) {
        const uint32_t s = fix_list[q];
        uint32_t* sb = d.st + st_index(s, 0);
        const bool search = wants_search(sb, head_all);             // CTA-uniform
        const bool first = blockIdx.x == 0 && threadIdx.x == 0;
        const uint32_t head = search_end(sb[F_POS * 32], head_all);
        if (search) {
            const uint32_t from = sb[F_POS * 32];
            const int k0 = (int) (sb[F_IDX * 32] >> 24);            // samples clocked since the last resume, before `from`
            const bool gap = k0 < kTapsUsed;                        // ... which is then at Rk = from - k0; older samples are the snapshot's
            const uint32_t Rk = from - (uint32_t) k0;
            const uint32_t base0 = from & ~7u;
            const uint32_t n_blk = (head - base0 + 7u) >> 3;
            for (uint32_t bi = blockIdx.x * kMaskThreads + threadIdx.x; bi < n_blk; bi += gridDim.x * kMaskThreads) {
                const uint32_t base = base0 + 8u * bi;
                Biquad b;
                if (bi == 0) { b.h1 = u2f(sb[F_IIR0 * 32]); b.h2 = u2f(sb[F_IIR1 * 32]); }
                else { const float2 c = d.ckpt_ring[blk_off(base - 8u, lmask, d.np, s)]; b.h1 = c.x; b.h2 = c.y; }
                // the filter outputs the block's correlations touch: 8 runs of 8 consecutive samples, 10 apart
                float w[8][8];
                if (!gap || (int32_t) (base - 70u - Rk) >= 0) {
#pragma unroll
                    for (int i = 0; i != 8; ++i) {
                        const uint32_t v = base - 70u + 10u * (uint32_t) i;        // even: the run is 4 aligned pairs
#pragma unroll
                        for (int j = 0; j != 8; j += 2) {
                            const float2 t2 = *reinterpret_cast<const float2*>(d.ring_y + ring_off(v + (uint32_t) j, lmask, d.np, s));
                            w[i][j] = t2.x; w[i][j + 1] = t2.y;
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i != 8; ++i)
#pragma unroll
                        for (int j = 0; j != 8; ++j) {
                            const uint32_t v = base - 70u + 10u * (uint32_t) i + (uint32_t) j;
                            float y;
                            if ((int32_t) (v - Rk) < 0) {
                                const int idx = 79 - (int) (Rk - 1u - v);
                                y = idx >= 0 ? d.corr_snap[(size_t) idx * d.np + s] : 0.f;   // idx < 0: older than 70 before `from`, never used
                            } else {
                                y = d.ring_y[ring_off(v, lmask, d.np, s)];
                            }
                            w[i][j] = y;
                        }
                }
                uint32_t word = 0;
#pragma unroll
                for (int j = 0; j != 8; ++j) {
                    const uint32_t p = base + (uint32_t) j;
                    if ((int32_t) (p - from) < 0 || (int32_t) (head - p) <= 0) continue;
                    const float limit = limit_step(b, w[7][j]);
                    const float y8[8] = {w[0][j], w[1][j], w[2][j], w[3][j], w[4][j], w[5][j], w[6][j], w[7][j]};
                    const uint32_t bits = trigger_bits(limit, corr8(y8, kWordPreamble), corr8(y8, kWordLsf), corr8(y8, kWordPacket), 0.f);
                    word |= ((bits & 1u) << j) | (((bits >> 1) & 1u) << (8 + j)) | (((bits >> 2) & 1u) << (16 + j));
                }
                d.mask_ring[blk_off(base, lmask, d.np, s)] = word;
            }
        }
        if (first && search) {
            sb[F_SFROM * 32] = sb[F_POS * 32]; sb[F_STO * 32] = head;
            atomicAdd(&d.ctr->search_requests, 1ull);
            atomicAdd(&d.ctr->search_samples, (unsigned long long) (head - sb[F_POS * 32]));
        }
    }
}

// The requests are served: the streams may be walked again.  Its own launch, behind the kernels above: the flags
// word is what hands a stream to whichever chain classifies it next, so everything else must be in place first.
__global__ void __launch_bounds__(128)
fixup_done_kernel(DevPtrs d, int cls)
{
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= d.ctr->cls[cls].fix_count) return;
    uint32_t* sb = d.st + st_index(d.fix_list[(size_t) cls * d.np + q], 0);
    sb[F_FLAGS * 32] &= ~FL_FIXUP;
}

void launch_fixup(const DevPtrs& d, uint32_t head, int sm_count, int cls, cudaStream_t s)
{
    fixup_kernel<<<sm_count * 8, kFixThreads, 0, s>>>(d, head, cls);
    search_limit_kernel<<<d.np / 128, 128, 0, s>>>(d, head, cls);
    // a request's span is at most kSearchSpan + 7 samples (one pass of x = its blocks), y = requests in strides of the grid
    search_mask_kernel<<<dim3(4, (unsigned) sm_count * 2), kMaskThreads, 0, s>>>(d, head, cls);
    fixup_done_kernel<<<d.np / 128, 128, 0, s>>>(d, cls);
}

void launch_decode(const DevPtrs& d, uint32_t head, unsigned long long head64, int cls, cudaStream_t s)
{
    size_t smem = sizeof(uint16_t) * 244 * kDecThreads;
    decode_kernel<<<(d.np + 5 * 32 + kDecThreads - 1) / kDecThreads, kDecThreads, smem, s>>>(d, head, head64, cls);
}

// Per-device kernel attributes (dynamic shared memory above 48 KiB).
cudaError_t configure_kernels()
{
    cudaError_t e = cudaFuncSetAttribute(decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int) (sizeof(uint16_t) * 244 * kDecThreads));
    if (e != cudaSuccess) return e;

    e = cudaFuncSetAttribute(fir_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int) (sizeof(float) * kFirStreams * kFirRow));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(fir_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int) (sizeof(float) * kFirStreams * kFirRow));
}

__global__ void init_state_kernel(DevPtrs d, uint32_t origin)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= d.np) return;
    init_lane(d.st + st_index(s, 0), origin);
}

void launch_init_state(const DevPtrs& d, uint32_t origin, cudaStream_t s)
{
    init_state_kernel<<<(d.np + 255) / 256, 256, 0, s>>>(d, origin);
}

// m17b_debug_correlate: Correlator::sample + correlate (Correlator.h:43-64) clocked on every sample of a fresh
// stream: limit_ and the correlations with the preamble, LSF / stream, packet / BERT and EOT sync words
// (M17Demodulator.h:154-157) per sample, from the matched-filter outputs in `d`'s ring.  One thread per stream
// (test-sized inputs); the arithmetic is limit_step / corr8, the very functions the search pass uses.
__global__ void debug_corr_kernel(DevPtrs d, uint32_t n, float* __restrict__ out /* [n_streams][n][5] */)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= d.n_streams) return;
    const uint32_t lmask = d.ring_lines - 1;
    Biquad b; b.h1 = 0.f; b.h2 = 0.f;
    for (uint32_t p = 0; p != n; ++p) {
        float y8[8];
#pragma unroll
        for (int i = 0; i != 8; ++i) {
            const int q = (int) p - 70 + 10 * i;
            y8[i] = q >= 0 ? d.ring_y[ring_off((uint32_t) q, lmask, d.np, s)] : 0.f;      // Correlator::buffer_ starts zeroed
        }
        float* o = out + ((size_t) s * n + p) * 5;
        o[0] = limit_step(b, y8[7]);
        o[1] = corr8(y8, kWordPreamble); o[2] = corr8(y8, kWordLsf); o[3] = corr8(y8, kWordPacket); o[4] = corr8(y8, kWordEot);
    }
}

void launch_debug_corr(const DevPtrs& d, uint32_t n, float* out, cudaStream_t s)
{
    debug_corr_kernel<<<(d.n_streams + 63) / 64, 64, 0, s>>>(d, n, out);
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
        } else if constexpr (VARIANT == 4) {
            // the exact-mode filter's instruction mix: two FMUL + one packed FADD2 per accumulator pair
            float mr = out[0] * 0.f + m;
            unsigned long long cc;
            asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
#pragma unroll
            for (int r = 0; r != 8; ++r)
#pragma unroll
                for (int i = 0; i != 16; i += 2) {
                    unsigned long long v;
                    float p0 = __fmul_rn(a[i], mr), p1 = __fmul_rn(a[i + 1], mr);
                    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(p0), "f"(p1));
                    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(v) : "l"(v), "l"(cc));
                    asm("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(a[i + 1]) : "l"(v));
                }
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
        case 4: fp32_peak_kernel<4><<<blocks, threads, 0, s>>>(buf, iters, 1.0f); break;
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
