/*
 * m17_port.h -- plain-C restatement ("port") of the reference's demodulation
 * path: M17Demodulator<float>::operator() and everything it calls
 * (/root/reference/include/m17cxx/M17Demodulator.h:656-753).
 *
 * TEST INFRASTRUCTURE ONLY (see m17_oracle.h).  Pinned against
 *   - the reference's own known-answer vectors (tests/test_oracle_kat.py), and
 *   - the unmodified reference compiled into oracle/_ref/libm17ref.so on
 *     identical inputs, record for record and trace entry for trace entry
 *     (tests/test_oracle_vs_ref.py).
 *
 * Differences from the reference that are deliberate:
 *   - the two function-local statics (`initializing` M17Demodulator.h:659,
 *     `eot_flag` :422) are per-demodulator members, i.e. every stream behaves
 *     like a fresh m17-demod process;
 *   - members the reference leaves indeterminate (output_buffer,
 *     depuncture_buffer; M17FrameDecoder.h:102-104) start as zero, which is what
 *     oracle/_ref gets too because it constructs the demodulator in calloc'd
 *     storage.
 */
#ifndef M17_PORT_H
#define M17_PORT_H

#include "m17_oracle.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct m17p_kalman { float x[2]; float P[2][2]; } m17p_kalman;

typedef struct m17p_syncword {
    int8_t  word[8];
    float   samples[10];
    float   mag1, mag2;
    uint8_t timing_index;
    uint8_t triggered;
    int8_t  updated;
} m17p_syncword;

typedef struct m17p_demod {
    /* BaseFirFilter<float,150> demod_filter  (FirFilter.h) */
    float    fir_hist[150];
    uint32_t fir_pos;
    /* DataCarrierDetect / NSlidingDFT<float,48000,120,2> */
    float    dft_samples[120];
    uint32_t dft_index;
    float    dft_re[2], dft_im[2];
    float    dcd_level_1, dcd_level_2, dcd_level;
    uint8_t  dcd_triggered;
    /* Correlator */
    float    corr_buf[80];
    float    corr_limit;
    uint32_t corr_pos, corr_prev_pos;
    float    iir_hist[3];
    m17p_syncword preamble_sync, lsf_sync, packet_sync, eot_sync;
    /* ClockRecovery */
    m17p_kalman clock_kf;
    uint64_t clock_count;
    int8_t   clock_sample_index;
    float    clock_estimate, sample_estimate;
    /* FreqDevEstimator */
    m17p_kalman dev_min, dev_max;
    float    dev_idev, dev_offset;
    uint8_t  dev_reset;
    /* SymbolEvm */
    float    evm_S;
    /* M17Framer<368> */
    int8_t   framer_buf[368];
    uint32_t framer_index;
    /* M17FrameDecoder */
    uint8_t  dec_state;
    uint8_t  dec_payload[26];     /* output_buffer union */
    uint8_t  dec_lsf[30];         /* output_buffer.lsf   */
    uint8_t  dec_out_type;
    uint8_t  dec_lich_segments;
    int8_t   dec_depunct[488];    /* depuncture_buffer union (stale tail matters for BERT) */
    /* M17Demodulator */
    uint64_t count;
    int8_t   polarity;
    uint8_t  demod_state;
    uint8_t  sync_word_type;
    uint8_t  sample_index;
    uint8_t  dcd;
    uint8_t  need_clock_reset, need_clock_update;
    uint64_t viterbi_cost;
    int32_t  sync_count, missing_sync_count;
    uint8_t  sync_sample_index;
    int16_t  initializing;
    uint8_t  eot_flag;
    /* output */
    m17o_record* rec;
    size_t   rec_cap, rec_n;
    uint32_t stream_id;
    uint64_t pos;
} m17p_demod;

void   m17p_init(m17p_demod* d);
void   m17p_sample(m17p_demod* d, float input);            /* M17Demodulator::operator() */
void   m17p_capture(const m17p_demod* d, m17o_trace* t);

/* Same contract as m17ref_demod_once, but re-entrant. */
size_t m17port_demod(const int16_t* samples, size_t n, unsigned flags, uint32_t stream_id,
                     m17o_record* rec, size_t rec_cap, m17o_trace* trace);
int    m17port_demod_batch(const int16_t* samples, size_t n_streams, size_t n_samples, size_t stride,
                           unsigned flags, m17o_record* rec, size_t rec_cap, uint64_t* counts,
                           m17o_trace* trace, int jobs_unused);

/* component taps mirroring the m17ref_* ones in ref_demod.cpp */
uint16_t m17port_crc16(const uint8_t* data, size_t n);
uint32_t m17port_golay_encode24(uint16_t data);
int      m17port_golay_decode(uint32_t input, uint32_t* output);
void     m17port_llr(float sample, int8_t* out2);
void     m17port_llr_table(float* thresholds43, int8_t* pairs86);
void     m17port_fir(const float* in, float* out, size_t n);
void     m17port_rrc_taps(float* taps150);
float    m17port_scale(int16_t s);
size_t   m17port_dcd(const float* in, size_t n, size_t period, float* level, uint8_t* trig);
void     m17port_sdft(const float* in, size_t n, float* out4);
void     m17port_limit_iir(const float* in, float* out, size_t n);
void     m17port_correlate_trace(const float* in, size_t n, float* out5);
void     m17port_correlator(const float* in, size_t n, uint8_t sample_index, float* corr4, float* mnmx2,
                            float* limit, uint8_t* index);
void     m17port_clock(const uint8_t* ops, const uint8_t* args, const uint32_t* dt, size_t n,
                       uint8_t* sample_index, float* clock_est, float* sample_est);
void     m17port_freqdev(const float* mn, const float* mx, const uint8_t* reset_before, size_t n,
                         float* offset, float* idev);
void     m17port_evm(const float* in, size_t n, float* out);
size_t   m17port_frame_decoder(const int8_t* frames, const uint8_t* sync, size_t n_frames,
                               m17o_record* rec, size_t rec_cap, int64_t* costs, uint8_t* states,
                               uint8_t* results);
size_t   m17port_viterbi(int kind, const int8_t* in, uint8_t* out);
void     m17port_tables(uint16_t* interleave368, int8_t* dc368, int8_t* p1_61, int8_t* p2_12, int8_t* p3_8);
void     m17port_sdft_coeffs(float* re_im4);

#ifdef __cplusplus
}
#endif
#endif
