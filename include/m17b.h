/*
 * m17b.h -- C ABI of libm17b.so, the B200-native many-channel M17 demodulator.
 *
 * The reference (mobilinkd/m17-cxx-demod) is a header-only C++ library with no C
 * ABI and no plugin mechanism; its demodulation entry points are
 *
 *   M17Demodulator<float>::M17Demodulator(callback_t)      M17Demodulator.h:180-182
 *   M17Demodulator<float>::operator()(float sample)        M17Demodulator.h:216,656-753
 *   M17Demodulator<float>::diagnostics(cb) / locked()      M17Demodulator.h:198-212
 *   M17FrameDecoder::callback_t(output_buffer_t const&, int viterbi_cost)
 *                                                          M17FrameDecoder.h:68-77,98
 *   the m17-demod stdin loop `demod(sample / 41067.0)`     apps/m17-demod.cpp:484-490
 *
 * This ABI is what a binding for that path binds instead: one context owns N
 * independent 48 kS/s int16 baseband streams resident on one GPU;
 * m17b_submit() replaces N x samples calls of operator()(float), and
 * m17b_drain() returns, per stream in chronological order, one m17b_frame per
 * invocation the reference would have made of callback_t -- bit for bit the
 * same output_buffer_t contents and viterbi cost.  The header-only facade
 * include/m17cxx_b200/M17Demodulator.h re-creates the reference's class API on
 * top of these calls.
 *
 * Conventions: plain C types; caller-allocated outputs; return 0 on success or
 * a negative M17B_E* code (m17b_strerror()); no exceptions cross the ABI; calls
 * on one context must be externally serialised; a context owns all of its
 * device memory and CUDA streams.  There is no CPU fallback: without a CUDA
 * device m17b_create() fails with M17B_ENODEV.
 */
#ifndef M17B_H
#define M17B_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define M17B_VERSION 0x00010000u

/* error codes */
#define M17B_OK        0
#define M17B_EINVAL   -1   /* bad argument                                         */
#define M17B_ENODEV   -2   /* no usable CUDA device / driver                       */
#define M17B_ENOMEM   -3   /* device or host allocation failed                     */
#define M17B_ECUDA    -4   /* a CUDA call or kernel failed (m17b_last_cuda_error)  */
#define M17B_EOVERFLOW -5  /* frame record ring overflowed: records were lost (counter `dropped`); returned once
                              per loss, the surviving records stay drainable: drain more often / raise
                              frame_capacity                                       */
#define M17B_ESTATE   -6   /* internal invariant violated (counter `faults`)       */
#define M17B_ETOOBIG  -7   /* size beyond the fixed scratch of m17b_filter_noalloc (M17B_NOALLOC_MAX_*) */

/* m17b_config.flags */
#define M17B_FIR_FMA     0x1u  /* matched filter accumulates with fused multiply-add: outputs within
                                  1e-5 relative of the reference but not bit-identical.  Default
                                  (flag clear) is the reference's exact mul-then-add order
                                  (FirFilter.h:36-40), bit-identical filter outputs.            */
#define M17B_INVERT      0x2u  /* `sample *= -1` before scaling (apps/m17-demod.cpp:488, -i)   */
#define M17B_PROFILE     0x4u  /* record CUDA events around every kernel (m17b_get_profile)    */
#define M17B_STREAMING   0x8u  /* back-to-back submits (m17b_ticket / m17b_drain_upto): between submits a
                                * stream may trail the newest sample, so that it always walks whole frames.
                                * A stream that is receiving delivers every frame that ends inside a submit
                                * with that submit; one that is acquiring (idle, searching for sync words, in
                                * a preamble) is served at its own pace and may deliver up to two submits
                                * (of max_chunk samples) later; the sample ring holds four.  m17b_sync / m17b_drain bring all streams up to the newest sample
                                * first (their diagnostics then match the reference's).  */

#define M17B_DIAG_RECORDS 0x10u /* record one m17b_diag_rec per reference diagnostic callback (m17b_diag_drain);
                                 * for few-channel callers such as the CLI: 56 bytes per 384 / 960 samples and stream */

/* m17b_submit location */
#define M17B_HOST    0         /* samples points to host memory (pinned for full overlap)      */
#define M17B_DEVICE  1         /* samples points to device memory on the context's GPU         */

/* frame types == M17FrameDecoder::FrameType (M17FrameDecoder.h:53) */
enum {
    M17B_FRAME_LSF = 0, M17B_FRAME_LICH = 1, M17B_FRAME_STREAM = 2,
    M17B_FRAME_BASIC_PACKET = 3, M17B_FRAME_FULL_PACKET = 4, M17B_FRAME_BERT = 5
};

typedef struct m17b_ctx m17b_ctx;

typedef struct m17b_config {
    uint32_t n_streams;        /* independent channels owned by this context               */
    int32_t  device;           /* CUDA device ordinal                                      */
    uint32_t max_chunk;        /* largest samples-per-stream a single m17b_submit may carry */
    uint32_t flags;            /* M17B_FIR_FMA | M17B_INVERT | M17B_PROFILE                */
    uint32_t frame_capacity;   /* frame records buffered on the device between drains
                                  (0: n_streams * 2 * (max_chunk / 1600 + 2), the worst case)  */
    uint32_t reserved[3];      /* zero.  reserved[0] is a test hook: ring position of every stream's first sample
                                  (32-bit ring positions wrap after 2^32 samples per stream; tests start near the wrap) */
} m17b_config;

/*
 * One record per reference callback invocation: output_buffer_t
 * (M17FrameDecoder.h:68-77) + the int viterbi_cost argument.  `payload` is the
 * 26-byte union {lich[6], stream[18], packet[26], bert[25]} exactly as the
 * reference's callback sees it (bytes past the active member are whatever an
 * earlier frame left there, like in the reference); `lsf` is output_buffer.lsf.
 */
typedef struct m17b_frame {
    uint32_t stream;           /* channel index in [0, n_streams)                          */
    uint32_t seq;              /* ordinal of this callback within its stream               */
    uint64_t sample_pos;       /* 0-based index of the input sample that fired it          */
    int32_t  viterbi_cost;
    uint8_t  type;             /* M17B_FRAME_*                                             */
    uint8_t  pad[3];
    uint8_t  payload[26];
    uint8_t  lsf[30];
} m17b_frame;                  /* 80 bytes */

/* The arguments of the reference's diagnostic_callback_t (M17Demodulator.h:144,
 * 681-685,746-750), sampled for one stream at its current position. */
typedef struct m17b_diag {
    int32_t  dcd;
    float    evm;
    float    deviation;
    float    offset;
    int32_t  locked;
    float    clock;
    int32_t  sample_index;
    int32_t  sync_index;
    int32_t  clock_index;
    int32_t  viterbi_cost;
    uint64_t samples;          /* samples of this stream processed so far                  */
} m17b_diag;

/* One record per invocation the reference would have made of diagnostic_callback_t: every 384 samples
 * while the carrier detect is off (evm argument = dev.error() = 0, M17Demodulator.h:675-689,
 * FreqDevEstimator.h:53) and every 960 samples while it is on (:742-752).  d.samples = samples of the
 * stream processed when it fires (the callback runs after the frame callback of the same sample).
 * Recorded only by contexts created with M17B_DIAG_RECORDS. */
typedef struct m17b_diag_rec {
    uint32_t  stream;
    uint32_t  pad;
    m17b_diag d;
} m17b_diag_rec;               /* 56 bytes */

typedef struct m17b_counters {
    uint64_t samples;          /* samples demodulated over all streams                     */
    uint64_t frames[6];        /* callbacks by M17B_FRAME_* type                           */
    uint64_t lsf_crc_fail;     /* LSF frames whose CRC16 failed (M17FrameDecoder.h:175)    */
    uint64_t golay_fail;       /* LICH frames with an uncorrectable Golay word (:195)      */
    uint64_t faults;           /* internal invariant violations (must stay 0)              */
    uint64_t dropped;          /* frame records lost to ring overflow (must stay 0)        */
    /* consumers of the frame records, per stream on the device (apps/m17-demod.cpp:207-253,287-305): */
    uint64_t packets_ok;       /* packets reassembled whose CRC-16/X.25 residue is 0x0f47  */
    uint64_t packets_crc_fail; /* last segment seen, checksum wrong                        */
    uint64_t packet_seq_errors;/* segments dropped: frame number out of sequence           */
    uint64_t packets_dropped;  /* packet records lost to ring overflow (must stay 0)       */
} m17b_counters;

/* What the reference CLI keeps per process on top of the frame callbacks, kept here per stream on the device
 * (decode_kernel feeds it): packet reassembly with sequence and CRC check (apps/m17-demod.cpp:80-175,207-253:
 * an LSF record clears the buffer -- and seeds it with the LSF for packet-mode types other than 1 --, every
 * non-final segment whose frame number is the expected one appends 25 bytes, the final segment appends its byte
 * count and the CRC-16/X.25 residue is checked; like the reference the buffer is not cleared by a final segment)
 * and PRBS9 bit-error accounting over BERT frames (decode_bert :287-305, Util.h:320-413, reset by the diagnostic
 * callback when it reports no carrier, :353-355). */
#define M17B_PACKET_MAX 860    /* 30 (LSF) + 33 segments of 25 bytes; a buffer that would grow past this stops growing */
typedef struct m17b_packet {
    uint32_t stream;
    uint32_t seq;              /* ordinal of this record within its stream                 */
    uint64_t sample_pos;       /* input sample that completed the final segment            */
    uint16_t length;           /* bytes in data[] (buffer contents incl. the 2 FCS bytes)  */
    uint8_t  fcs_ok;           /* CRC-16/X.25 residue == 0x0f47                            */
    uint8_t  pad;
    uint8_t  data[M17B_PACKET_MAX];
} m17b_packet;                 /* 880 bytes */

typedef struct m17b_ber {
    int32_t  synced;           /* PRBS9::synced                                            */
    uint32_t bits;             /* PRBS9::bit_count                                         */
    uint32_t errors;           /* PRBS9::err_count                                         */
    uint32_t packet_bytes;     /* bytes currently in the stream's packet buffer            */
} m17b_ber;

/* Accumulated device time per kernel (CUDA events on the launching stream),
 * only when M17B_PROFILE is set. */
typedef struct m17b_profile {
    double   fir_ms;     uint64_t fir_launches;     uint64_t fir_samples;    /* matched filter: outputs computed */
    double   scan_ms;    uint64_t scan_launches;    uint64_t scan_samples;   /* sync/clock/slicer scan           */
    double   decode_ms;  uint64_t decode_launches;  uint64_t decode_frames;  /* Viterbi/Golay/CRC frame decode   */
    double   aux_ms;     uint64_t aux_launches;                              /* history tail upkeep              */
    double   h2d_ms;     uint64_t h2d_bytes;
    double   d2h_ms;     uint64_t d2h_bytes;
    double   dcd_ms;     uint64_t dcd_launches;                              /* carrier-detect pre-pass          */
    double   fixup_ms;   uint64_t fixup_launches;                            /* filter fix-up + sync-word search */
    uint64_t search_requests; uint64_t search_samples;                       /* searches served / samples covered */
    double   scan1_ms;                                                       /* scan of the acquiring streams (walk class 1; scan_ms = receiving streams) */
} m17b_profile;

uint32_t    m17b_version(void);
const char* m17b_strerror(int code);
const char* m17b_last_cuda_error(const m17b_ctx* ctx);

int  m17b_device_count(void);

int  m17b_create(const m17b_config* cfg, m17b_ctx** out);
void m17b_destroy(m17b_ctx* ctx);

/* Pinned host memory helpers (optional; any host pointer works with M17B_HOST). */
int  m17b_host_alloc(void** ptr, size_t bytes);
void m17b_host_free(void* ptr);
/* Device memory on the context's GPU, for callers that stage samples themselves. */
int  m17b_device_alloc(m17b_ctx* ctx, void** ptr, size_t bytes);
void m17b_device_free(m17b_ctx* ctx, void* ptr);
int  m17b_memcpy_h2d(m17b_ctx* ctx, void* dst_device, const void* src_host, size_t bytes);

/*
 * Feed `samples_per_stream` new int16 samples to every stream: stream s reads
 * samples[s * stream_stride + 0 .. samples_per_stream).  Asynchronous: returns
 * once the work is enqueued; the buffer must stay valid until the next
 * m17b_sync()/m17b_drain().  samples_per_stream <= cfg.max_chunk.
 */
int  m17b_submit(m17b_ctx* ctx, const int16_t* samples, size_t samples_per_stream,
                 size_t stream_stride, int location);

/* Wait for all enqueued work; every stream has then consumed all submitted samples. */
int  m17b_sync(m17b_ctx* ctx);

/* m17b_sync(), then move up to `cap` frame records to `out`.  Records of one stream appear in
 * chronological (seq) order; records of different streams are interleaved in the order the device
 * completed them.  *n receives the number written; call again while *n == cap. */
int  m17b_drain(m17b_ctx* ctx, m17b_frame* out, size_t cap, size_t* n);

/* Pipelined draining.  Every m17b_submit() gets a ticket (1, 2, ...; m17b_ticket() returns the last one
 * issued).  m17b_drain_upto() waits only until the submit with that ticket has finished on the device --
 * later submits keep running -- and returns the records appended up to then (same ordering rules as
 * m17b_drain; a stream that had not reached the end of that submit yet delivers the rest with a later
 * ticket or with m17b_drain).  The device-to-host transfer runs on the copy stream beside the kernels of
 * the later submits.  Keep at most 8 submits in flight -- and size m17b_config.frame_capacity for the depth you use: the
 * default holds the worst case of three submits (the one being drained and two enqueued behind it).  When the ring does
 * overflow, the surviving records stay drainable and M17B_EOVERFLOW is returned once per loss (counter `dropped`). */
unsigned long long m17b_ticket(const m17b_ctx* ctx);
int  m17b_drain_upto(m17b_ctx* ctx, unsigned long long ticket, m17b_frame* out, size_t cap, size_t* n);

/* Number of frame records currently waiting on the device (after m17b_sync). */
int  m17b_pending_frames(m17b_ctx* ctx, size_t* n);

int  m17b_diag_get(m17b_ctx* ctx, uint32_t stream, m17b_diag* out);
/* m17b_sync(), then move up to `cap` diagnostic records (M17B_DIAG_RECORDS contexts only) to `out`, per stream in
 * chronological order; call again while *n == cap. */
int  m17b_diag_drain(m17b_ctx* ctx, m17b_diag_rec* out, size_t cap, size_t* n);
int  m17b_counters_get(m17b_ctx* ctx, m17b_counters* out);
/* m17b_sync(), then move up to `cap` packet records (one per final packet segment a stream decoded, checksum right or
 * wrong) to `out`, per stream in chronological order; call again while *n == cap. */
int  m17b_packet_drain(m17b_ctx* ctx, m17b_packet* out, size_t cap, size_t* n);
/* The stream's PRBS9 accounting (what the reference prints as "BER: errors / bits") after m17b_sync(). */
int  m17b_ber_get(m17b_ctx* ctx, uint32_t stream, m17b_ber* out);
int  m17b_get_profile(m17b_ctx* ctx, m17b_profile* out, int reset);
/* Switch the per-kernel event records of M17B_PROFILE on or off for the submits that follow (m17b_sync first), so that a
 * benchmark can time its steps without them and take the kernel shares from a separate pass. */
int  m17b_profile_enable(m17b_ctx* ctx, int on);

/* Device-side timing for benchmarks: record CUDA event `slot` (0..7) on the context's compute
 * stream now; elapsed milliseconds between two recorded slots (synchronises on `b`). */
int  m17b_event_record(m17b_ctx* ctx, int slot);
int  m17b_event_elapsed(m17b_ctx* ctx, int a, int b, double* ms);

/* Kernel-level taps for parity tests: run ONLY the matched filter over
 * [n_streams][n] int16 (zero history) and return the n filtered floats per stream,
 * in the filter mode the context was created with. */
int  m17b_debug_filter(m17b_ctx* ctx, const int16_t* samples_host, size_t n, size_t stream_stride,
                       float* filtered_host /* [n_streams][n] */, float* scaled_host /* may be NULL */);

/* ... and Correlator::sample + correlate (Correlator.h:43-64) clocked on every one of those filter outputs from a
 * fresh correlator: per sample {limit_, correlation with the preamble, LSF/stream, packet/BERT, EOT sync word}
 * (M17Demodulator.h:154-157), computed by the arithmetic the sync-word search pass uses. */
int  m17b_debug_correlate(m17b_ctx* ctx, const int16_t* samples_host, size_t n, size_t stream_stride,
                          float* filtered_host /* [n_streams][n] */, float* corr_host /* [n_streams][n][5] */);

/*
 * No counterpart in the reference.  The matched filter of
 * m17b_debug_filter (scaling + 150-tap RRC, zero history, filter mode from `flags`: 0, M17B_FIR_FMA, M17B_INVERT) for
 * callers that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
 * It calls no cudaMalloc / cudaFree / allocating routine and needs no m17b_ctx: the kernel's scratch (its temporary
 * sample ring, 3,186,688 bytes) is a static __device__ array that every GPU which loads the library holds until exit.
 * Hence the limits: n_streams <= M17B_NOALLOC_MAX_STREAMS and n <= M17B_NOALLOC_MAX_SAMPLES, beyond which it returns
 * M17B_ETOOBIG (use m17b_debug_filter).  Inputs and outputs are the caller's DEVICE buffers on `device`:
 * samples_device[s * stream_stride + 0 .. n) int16, filtered_device [n_streams][n] float.  The work is enqueued on
 * `cuda_stream` (a cudaStream_t; NULL = the default stream) and the call returns without waiting for it.  All calls on
 * one GPU share the scratch array (it keeps the last call's contents; what must start at zero is cleared in every
 * call): calls that overlap on the same GPU race -- serialise them, e.g. by using one stream.
 */
#define M17B_NOALLOC_MAX_STREAMS 128u
#define M17B_NOALLOC_MAX_SAMPLES 4000u
int  m17b_filter_noalloc(int device, const int16_t* samples_device, uint32_t n_streams, size_t n, size_t stream_stride,
                         float* filtered_device, uint32_t flags, void* cuda_stream);

/* FP32 pipe microbenchmark: sustained FFMA throughput of the context's GPU in TFLOP/s
 * (2 flop per FMA), the denominator of the matched filter's roofline fraction.
 * variant 0: scalar FFMA; 1: packed fma.rn.f32x2; 2: FMUL+FADD pairs;
 * 3: register-operand FFMA; 4: two FMUL + one packed FADD2 (the exact-mode filter mix). */
int  m17b_measure_fp32_peak(m17b_ctx* ctx, int variant, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
