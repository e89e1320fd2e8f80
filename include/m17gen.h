/*
 * m17gen.h -- C ABI of the synthetic M17 baseband generator (libm17gen.so).
 *
 * The generator is the input side of every parity test and of bench.py: it
 * restates the frame construction of the reference's transmit CLI
 * (/root/reference/apps/m17-mod.cpp, the USE_OLD_MODULATOR path that m17-mod
 * actually ships, :204-564) with random payloads instead of Codec2 audio, adds
 * packet-mode frames (which m17-mod cannot emit) built from the decoder's own
 * definition (M17FrameDecoder.h:299-315, Trellis.h:39-40, sync word
 * M17Demodulator.h:156), and applies seeded channel impairments (AWGN, symbol
 * clock skew, DC offset == the demodulator's "frequency offset").
 *
 * libm17gen.so is CPU code; libm17gen_cuda.so (m17gen_batch_device) builds the same waveforms on the GPU.
 * Neither is on the demodulation path.
 */
#ifndef M17GEN_H
#define M17GEN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    M17GEN_STREAM = 0,  /* preamble, LSF, n stream frames (last has EOS), EOT   */
    M17GEN_PACKET = 1,  /* preamble, LSF(type packet), n packet frames, EOT     */
    M17GEN_BERT   = 2,  /* preamble, LSF(type BERT), n BERT frames, EOT         */
    M17GEN_MIXED  = 3   /* a seeded sequence of the three kinds + short streams */
};

typedef struct m17gen_params {
    uint64_t seed;
    int32_t  mode;            /* M17GEN_*                                          */
    int32_t  n_frames;        /* payload frames per transmission (<=0: fill buffer)*/
    int32_t  lead_in;         /* samples of channel noise before the preamble      */
    int32_t  tail;            /* minimum samples of noise after EOT                */
    int32_t  packet_encapsulated; /* packet LSF: 0 = RAW (type 1), 1 = ENCAPSULATED (2) */
    int32_t  invert;          /* multiply the waveform by -1                       */
    int32_t  repeat;          /* 1: start a new transmission after `gap` samples until the buffer is full */
    int32_t  gap;             /* idle samples between repeated transmissions       */
    float    snr_db;          /* sample-domain SNR of the AWGN vs the modulated signal's RMS; >= 200: no noise */
    float    skew_ppm;        /* transmitter symbol clock error, +: TX clock fast  */
    float    dc_offset;       /* additive offset in int16 counts                   */
    float    gain;            /* amplitude scale on top of the reference's 7168    */
    float    phase;           /* initial fractional sample phase in [0,10); <0: seeded random */
    float    idle_noise_rms;  /* RMS (counts) of the noise floor when snr_db >= 200 or during idle; 0 => exact zeros */
} m17gen_params;

/* What was transmitted, for BER sanity checks (not used for parity). */
typedef struct m17gen_truth {
    uint32_t n_transmissions;
    uint32_t n_stream_frames;
    uint32_t n_packet_frames;
    uint32_t n_bert_frames;
    uint8_t  first_lsf[30];
    uint8_t  pad[2];
} m17gen_truth;

void m17gen_default_params(m17gen_params* p);

/* Generate one stream of n_samples int16 (48 kS/s).  Returns 0 on success. */
int m17gen_stream(const m17gen_params* p, int16_t* out, size_t n_samples, m17gen_truth* truth);

/* n_streams streams, params[i] each, out[i*stride ...]; n_threads host threads. */
int m17gen_batch(const m17gen_params* params, size_t n_streams, int16_t* out, size_t n_samples,
                 size_t stride, m17gen_truth* truth /* may be NULL */, int n_threads);

/* 10^(snr_db / 20) as the generator computes it (host libm; the device-side build takes it as an input). */
double m17gen_snr_linear(float snr_db);

/* Device-side modulator (libm17gen_cuda.so, csrc/m17gen_dev.cu): the same n_streams streams written to device
 * memory out_device[i * stride ...] on CUDA device `device`, bit-identical to m17gen_batch for the same
 * parameters.  params and snr_linear (one m17gen_snr_linear(params[i].snr_db) per stream) are host arrays.
 * Returns 0 on success. */
int m17gen_batch_device(const m17gen_params* params, const double* snr_linear, size_t n_streams,
                        int16_t* out_device, size_t n_samples, size_t stride, int device);

/* Building blocks, exposed so tests can pin them against the reference's KATs. */
uint16_t m17gen_crc16(const uint8_t* data, size_t n);               /* CRC16.h, poly 0x5935 */
uint32_t m17gen_golay24_encode(uint16_t data12);                    /* Golay24.h:184-201   */
void     m17gen_lsf_frame_bits(const uint8_t lsf[30], uint8_t out368[368]);   /* m17-mod.cpp:354-384 */
void     m17gen_stream_frame_bits(const uint8_t lsf[30], int lich_index, uint16_t fn,
                                  const uint8_t payload16[16], uint8_t out368[368]); /* :407-440,:509-564 */
void     m17gen_packet_frame_bits(const uint8_t data26[26], uint8_t out368[368]);
void     m17gen_bert_frame_bits(uint16_t* prbs_state, uint8_t out368[368]);   /* :442-504 */
void     m17gen_encode_callsign(const char* call, uint8_t out6[6]); /* LinkSetupFrame.h:48-92 */

#ifdef __cplusplus
}
#endif
#endif
