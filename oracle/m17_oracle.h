/*
 * m17_oracle.h -- shared record / trace layouts for the CPU oracles.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load or execute it, and only as the checker or
 * as the timed CPU baseline.  The product (m17_cxx_demod_b200/) never links,
 * imports or falls back to anything in this directory.
 *
 * Two implementations produce these records from identical int16 baseband:
 *   oracle/_ref/libm17ref.so  -- the UNMODIFIED reference headers
 *                                (/root/reference/include/m17cxx/M17Demodulator.h)
 *                                compiled by oracle/Makefile  ("reference")
 *   oracle/libm17port.so      -- oracle/m17_port.c, a plain-C restatement
 *                                of the same algorithm                ("port")
 * The product's CUDA path must reproduce the same records bit for bit.
 */
#ifndef M17_ORACLE_H
#define M17_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Frame types: M17FrameDecoder::FrameType (M17FrameDecoder.h:53). */
enum {
    M17O_LSF = 0, M17O_LICH = 1, M17O_STREAM = 2,
    M17O_BASIC_PACKET = 3, M17O_FULL_PACKET = 4, M17O_BERT = 5
};

/*
 * One record per invocation of the reference's frame callback
 * (M17FrameDecoder::callback_t, M17FrameDecoder.h:98): the complete
 * output_buffer_t (M17FrameDecoder.h:68-77) plus the viterbi cost argument.
 * `payload` is the 26-byte union {lich[6], stream[18], packet[26], bert[25]}
 * exactly as the callback sees it (stale bytes beyond the active member
 * included); `lsf` is output_buffer.lsf.
 */
typedef struct m17o_record {
    uint32_t stream;       /* stream index inside the batch                 */
    uint32_t seq;          /* ordinal of this callback within the stream    */
    uint64_t sample_pos;   /* 0-based index of the input sample whose
                              processing fired the callback                 */
    int32_t  viterbi_cost; /* the callback's int argument                   */
    uint8_t  type;         /* M17O_* frame type                             */
    uint8_t  pad[3];
    uint8_t  payload[26];
    uint8_t  lsf[30];
} m17o_record;             /* 80 bytes */

/*
 * Per-sample trace of demodulator internals (all public members of
 * M17Demodulator, M17Demodulator.h:148-178), captured AFTER operator() has
 * processed the sample.  Used to localise the first divergence between two
 * implementations.
 */
typedef struct m17o_trace {
    float    filtered;      /* correlator.buffer_[prev_buffer_pos_]: last FIR output clocked in */
    float    limit;         /* correlator.limit()              */
    float    dcd_level;     /* dcd.level()                     */
    float    dev_offset;    /* dev.offset()                    */
    float    dev_idev;      /* dev.idev()                      */
    float    clock_est;     /* clock_recovery.clock_estimate() */
    float    sample_est;    /* clock_recovery.sample_estimate_ */
    float    evm;           /* evm.evm()                       */
    int32_t  sync_count;
    int32_t  missing_sync_count;
    int32_t  viterbi_cost;  /* (int) viterbi_cost              */
    uint32_t count;         /* (uint32) count_                 */
    uint8_t  demod_state;   /* DemodState                      */
    uint8_t  sample_index;
    uint8_t  sync_sample_index;
    uint8_t  clock_sample_index; /* clock_recovery.sample_index() */
    uint8_t  dcd;           /* dcd_                            */
    uint8_t  dcd_triggered; /* dcd.dcd()                       */
    uint8_t  sync_word_type;
    uint8_t  decoder_state; /* decoder.state()                 */
    uint8_t  corr_index;    /* correlator.index()              */
    uint8_t  need_clock_reset;
    uint8_t  need_clock_update;
    uint8_t  polarity;
} m17o_trace;               /* 60 bytes */

/* flags */
#define M17O_INVERT      1u  /* apps/m17-demod.cpp:488 `sample *= -1`          */
#define M17O_EOF_REPEAT  2u  /* replay the CLI's `while (std::cin)` quirk:
                                the last sample is processed twice
                                (apps/m17-demod.cpp:484-490)                   */

/*
 * Demodulate ONE stream.  `rec` receives up to `rec_cap` records (extra
 * callbacks are counted but not stored); `trace`, when non-NULL, must hold n
 * (+1 with M17O_EOF_REPEAT) entries.  Returns the number of callbacks fired.
 */
typedef size_t (*m17o_demod_fn)(const int16_t* samples, size_t n, unsigned flags,
                                uint32_t stream_id, m17o_record* rec, size_t rec_cap,
                                m17o_trace* trace);

#ifdef __cplusplus
}
#endif
#endif
