// m17b_dev.cuh -- device-side layout shared by the kernels and the host context.
//
// Data layout in HBM (N = streams padded to a multiple of 128, "Np"):
//
//   sample ring   ring_y : float [RL lines][Np][32]
//       line l of stream s holds samples 32l .. 32l+31 (128 B = one L2 line, four
//       32-byte sectors): the matched filter output computed speculatively
//       for every sample.  The lanes of a scan warp sit at different
//       stream positions (each stops at its own frame completions), so a lane must
//       own whole lines: it consumes the four sectors of a line in four consecutive
//       8-sample blocks (one DRAM fetch, three L2 hits).  With 8-sample lines
//       interleaved over streams (the first layout) every sector read dragged in
//       three neighbours' sectors that were evicted before their lanes got there:
//       3.6x the algorithmic DRAM traffic (profiles/r01b_*).  (Round 1 kept a second such
//       ring for x[n] - x[n-120], the carrier detect's input; dcd_kernel now forms it
//       from the int16 input.)
//                 ring_x : int16 [RL][Np][32]  raw (sign-corrected) input, only for the
//       rare paths that rebuild filter history after a carrier-detect gap.
//   per-stream scalars            : uint32 [Np/32][NFIELDS][32]   (one coalesced
//       128 B line per field per warp)
//   sync-word trigger masks       : uint32 [RL][Np][4]   one word per 8-sample block: bit 8w + j set when sync word w
//                                   (0 preamble, 1 LSF/stream, 2 packet/BERT) triggers on block sample j
//                                   (Correlator.h:150-157); written by fixup_kernel's search pass for the span a
//                                   searching stream (ST_UNLOCKED, carrier detect on) has ahead of it
//   limit-filter checkpoints      : float2 [RL][Np][4]   history of Correlator::sample_filter after every block
//                                   of a searched span (the stream adopts it where it leaves the search)
//   correlator snapshot           : float [80][Np]    (Correlator::buffer_ frozen when the carrier detect gates it)
//   carrier-detect window sums    : float [n_seg][4][Np]  (dcd_kernel: energy sums of both sliding-DFT bins
//                                   384 and 960 samples after every 192-aligned window start)
//   sync-word peak samples        : float [30][Np]    (SyncWord::samples_ x3)
//   FIR history snapshot          : float [149][Np]   (taken when DCD gates the filter)
//   soft-bit frame                : int16 [184][Np]   (M17Framer buffer, 2 LLRs per symbol)
//   decoder bytes                 : uint8 [56][Np]    (output_buffer union[26] + lsf[30])
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "m17b.h"

namespace m17b {

constexpr int kTaps = 150;            // M17Demodulator.h:79-118 (tap 149 is 0.0)
constexpr int kTapsUsed = 149;
constexpr int kTailLen = 160;         // raw history kept between submits (>= 148 + 7)
constexpr int kChunk = 8;             // samples per scan block (one 32-byte sector)
constexpr int kLine = 32;             // samples per ring line (128 bytes of float)
constexpr int kDftN = 120;            // NSlidingDFT length (M17Demodulator.h:149: 48000/400)
constexpr int kRoundCap = 2048;       // ring slack: samples a stream may trail the head by beyond one submit
constexpr int kMinFrameSpacing = 1600;// lower bound on samples between two frame completions
constexpr int kSeg = 192;             // gcd of the carrier-detect cadences 384 / 960 and of the 1920-sample start-up
constexpr int kPktStride = 864;       // bytes of packet buffer per stream (>= M17B_PACKET_MAX, multiple of 16)
constexpr int kClasses = 2;           // walk classes a round sorts the streams into (classify_kernel)
constexpr int kFirstUpdate = 1920 + 384;   // position of a stream's first carrier-detect update (M17Demodulator.h:659-689)

enum Field : int {
    F_POS = 0,
    F_RE0, F_IM0, F_RE1, F_IM1, F_L1, F_L2, F_LEVEL,
    F_IIR0, F_IIR1, // limit biquad history (Correlator.h:38-41) at F_POS; frozen while the carrier detect gates it
    F_SFROM, F_STO, // span [from, to) the search pass has covered for this stream's current search
    F_CNT,          // count_ (low 16) | initializing (high 16)
    F_IDX,          // corr_pos (8) | framer_index (16) << 8 | gap_k (8) << 24
    F_FLAGS,
    F_SW,           // SyncWord timing_index/triggered/updated x3
    F_SYNC_COUNT, F_MISSING, F_VCOST,
    F_CK_COUNT, F_CK_EST, F_CK_SEST,
    F_CK_X0, F_CK_X1, F_CK_P00, F_CK_P01, F_CK_P10, F_CK_P11,
    F_DMIN_X0, F_DMIN_X1, F_DMIN_P00, F_DMIN_P01, F_DMIN_P10, F_DMIN_P11,
    F_DMAX_X0, F_DMAX_X1, F_DMAX_P00, F_DMAX_P01, F_DMAX_P10, F_DMAX_P11,
    F_IDEV, F_OFFSET, F_EVM,
    F_DEC,          // dec_state (8) | lich_segments (8) << 8 | depunct401 (8) << 16
    F_SEQ,
    F_GAPR,         // position at which the matched filter last resumed after a carrier-detect gap
    F_GAPDONE,      // outputs after F_GAPR already corrected by fixup_kernel (0..149)
    F_ACC1,         // dcd_kernel: five rolling carrier-detect energy sums of bin 0 (window phases), F_ACC1 + i
    F_ACC2 = F_ACC1 + 5,   // ... of bin 1
    // consumers of the frame records (apps/m17-demod.cpp's per-process state, per stream here)
    F_PKT = F_ACC2 + 5,    // packet buffer length (16) | packet_frame_counter (8) << 16 | packet ordinal (8) << 24
    F_PRBS,         // PRBS9: state (9) | synced << 9 | sync_count (5) << 10 | hist_count (8) << 15 | hist_pos (7) << 23 | reset after the pending frame << 30
    F_BER_BITS, F_BER_ERRS,
    F_BER_H0,       // PRBS9::history, 128 bits, F_BER_H0 + i
    NFIELDS = F_BER_H0 + 4
};

// F_FLAGS bit layout
constexpr uint32_t FL_STATE_SHIFT = 0;    // 3 bits, DemodState; 7 = decided by the decode kernel
constexpr uint32_t FL_SWT_SHIFT = 3;      // 2 bits, SyncWordType
constexpr uint32_t FL_SI_SHIFT = 5;       // 4 bits, sample_index
constexpr uint32_t FL_SSI_SHIFT = 9;      // 4 bits, sync_sample_index
constexpr uint32_t FL_CSI_SHIFT = 13;     // 4 bits, clock_recovery.sample_index_
constexpr uint32_t FL_DCD = 1u << 17;
constexpr uint32_t FL_DCD_TRIG = 1u << 18;
constexpr uint32_t FL_NEED_RESET = 1u << 19;
constexpr uint32_t FL_NEED_UPDATE = 1u << 20;
constexpr uint32_t FL_DEV_RESET = 1u << 21;
constexpr uint32_t FL_EOT = 1u << 22;
constexpr uint32_t FL_PENDING = 1u << 23;     // a frame is complete: decode_kernel's turn
constexpr uint32_t FL_FIXUP = 1u << 24;       // filter outputs after a gap and / or a sync-word search are due: fixup_kernel's turn

enum DemodState : int { ST_UNLOCKED, ST_LSF_SYNC, ST_STREAM_SYNC, ST_PACKET_SYNC, ST_BERT_SYNC, ST_SYNC_WAIT, ST_FRAME, ST_TBD };
enum DecState : int { DEC_LSF, DEC_STREAM, DEC_BASIC_PACKET, DEC_FULL_PACKET, DEC_BERT };
enum SyncType : int { SW_LSF, SW_STREAM, SW_PACKET, SW_BERT };

struct ClassCounters {            // work lists of one walk class, reset before the class's chain of a round runs
    unsigned int scan_count;      // entries in scan_list
    unsigned int fix_count;       // entries in fix_list
    unsigned int dec_count[5];    // entries in dec_list per decode kind
    unsigned int lagging;         // streams of the class that have not reached the head
    unsigned int deferred;        // streams that left a partial frame for the next submit (M17B_STREAMING)
    unsigned int pad;
};

struct Counters {                 // device-side mirror of m17b_counters (+ record ring cursor)
    unsigned long long frames[6];
    unsigned long long lsf_crc_fail;
    unsigned long long golay_fail;
    unsigned long long faults;
    unsigned long long dropped;
    unsigned int rec_count;       // record cursor: records appended so far (monotonic, mod 2^32)
    unsigned int diag_count;      // diagnostic-record cursor (M17B_DIAG_RECORDS), monotonic like rec_count
    ClassCounters cls[kClasses];
    unsigned long long diag_dropped;
    unsigned long long search_requests;   // sync-word searches served by search_mask_kernel
    unsigned long long search_samples;    // ... and the samples they covered
    unsigned long long packets_ok, packets_crc_fail, packet_seq_errors, packets_dropped;
    unsigned int pkt_count;       // packet-record cursor, monotonic like rec_count
    unsigned int pad1;
};

struct DevPtrs {
    uint32_t* st;                 // [Np/32][NFIELDS][32]
    float* ring_y;                // [RL][Np][32]
    int16_t* ring_x;              // [RL][Np][32]
    uint32_t* mask_ring;          // [RL][Np][4] sync-word trigger masks, one word per 8-sample block
    float* corr_snap;             // [80][Np] the correlator's last 80 samples before a carrier-detect gap
    float* seg_ring;              // [n_seg][4][Np] carrier-detect energy sums per window start (see dcd_lane)
    float2* ckpt_ring;            // [RL][Np][4] limit-filter history after every 8-sample block of a searched span
    uint32_t* fix_list;           // [kClasses][Np] streams waiting for fixup_kernel, by the walk class that asked
    uint32_t* scan_list;          // [kClasses][Np] streams with samples to walk this round, by walk class (classify_kernel):
                                  // 0 = receiving (carrier detect on, in or between frames), 1 = everything else (idle,
                                  // start-up, sync-word search, preamble, after a carrier-detect gap)
    uint32_t* dec_list;           // [kClasses][5][Np] streams with a completed frame, by walk class and decode kind (0..3 = Viterbi map, 4 = LICH / none)
    float* sw_samples;            // [30][Np]
    float* fir_snap;              // [149][Np]
    int16_t* llr;                 // [184][Np]
    uint8_t* dec_bytes;           // [56][Np]
    m17b_frame* recs;             // [rec_cap]
    m17b_diag_rec* diag_recs;     // [diag_cap] or null: diagnostics at the reference's update points (M17B_DIAG_RECORDS)
    uint8_t* pkt_buf;             // [Np][kPktStride] packet reassembly buffer of every stream (current_packet)
    m17b_packet* pkt_recs;        // [pkt_cap]
    Counters* ctr;
    uint32_t n_streams;
    uint32_t np;                  // padded stream count (multiple of 128)
    uint32_t ring_lines;          // RL, power of two
    uint32_t n_seg;               // entries of seg_ring, power of two, n_seg * 192 > ring span + 960
    uint32_t rec_cap;             // power of two
    uint32_t rec_read;            // cursor up to which the host has taken the records (at launch time)
    uint32_t diag_cap;            // power of two
    uint32_t diag_read;
    uint32_t pkt_cap;             // power of two
    uint32_t pkt_read;
};

__host__ __device__ inline size_t st_index(uint32_t s, int f)
{
    return (size_t) (s >> 5) * (NFIELDS * 32) + (size_t) f * 32 + (s & 31);
}

// Viterbi input maps: one int16 per depunctured soft bit.  0 = erasure,
// +-(k+1) = +-frame[k] (derandomise sign folded in, deinterleave folded in),
// kMapStale = the stale depuncture_buffer element a BERT frame inherits.
constexpr int16_t kMapStale = 0x7FFF;
// kind: 0 LSF, 1 stream, 2 packet, 3 BERT (M17FrameDecoder.h:79-91)
__host__ __device__ constexpr int map_len(int kind) { return kind == 0 ? 488 : kind == 1 ? 296 : kind == 2 ? 420 : 402; }
__host__ __device__ constexpr int out_bits(int kind) { return kind == 0 ? 240 : kind == 1 ? 144 : kind == 2 ? 206 : 197; }

struct ConstTables {
    float taps[kTaps];
    float sdft[4];                // re0, im0, re1, im1
    float llr_thr[43];
    int8_t llr_val[43][2];
    int16_t vmap[4][488];
    int16_t lichmap[96];          // LICH soft bits 0..95 of the deinterleaved frame
    uint16_t crc_init;
};

// host side (m17b_tables.cpp)
void build_tables(ConstTables& t, uint32_t* golay2048);

// launchers (m17b_kernels.cu)
cudaError_t upload_tables(const ConstTables& t, const uint32_t* golay2048);
void launch_fir(const DevPtrs& d, const int16_t* in, size_t stride, uint32_t n_new, const int16_t* tail,
                uint32_t head, bool fma, bool invert, cudaStream_t s);
void launch_tail(const DevPtrs& d, const int16_t* in, size_t stride, uint32_t n_new, const int16_t* tail_in,
                 int16_t* tail_out, cudaStream_t s);
void launch_dcd(const DevPtrs& d, const int16_t* in, size_t stride, const int16_t* tail, uint32_t head_old, uint32_t n_new,
                unsigned long long head64_old, bool invert, cudaStream_t s);
void launch_classify(const DevPtrs& d, uint32_t head, uint32_t class_mask, bool defer, cudaStream_t s);
void launch_scan(const DevPtrs& d, uint32_t head, unsigned long long head64, bool defer, uint32_t guard, int cls, bool split, cudaStream_t s);
void launch_decode(const DevPtrs& d, uint32_t head, unsigned long long head64, int cls, cudaStream_t s);
void launch_fixup(const DevPtrs& d, uint32_t head, int sm_count, int cls, cudaStream_t s);
void launch_gather(const DevPtrs& d, uint32_t n, float* y, float* x, cudaStream_t s);
void launch_init_state(const DevPtrs& d, uint32_t origin, cudaStream_t s);
void launch_debug_corr(const DevPtrs& d, uint32_t n, float* out, cudaStream_t s);
cudaError_t configure_kernels();
double run_fp32_peak(int variant, int sm_count, cudaStream_t s);

} // namespace m17b
