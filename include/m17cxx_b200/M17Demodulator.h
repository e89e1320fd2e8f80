// M17Demodulator.h -- the reference's demodulator class API on top of libm17b.so.
//
// Mirrors, for the demodulation path only, what an application includes from the
// reference (include/m17cxx/M17Demodulator.h + M17FrameDecoder.h):
//
//   mobilinkd::M17FrameDecoder::{FrameType, output_buffer_t, callback_t}
//                                              M17FrameDecoder.h:50-53,68-77,98
//   mobilinkd::M17Demodulator<FloatType>
//       M17Demodulator(callback_t)             M17Demodulator.h:180-182
//       void operator()(FloatType sample)      M17Demodulator.h:216
//       void diagnostics(diagnostic_callback_t)M17Demodulator.h:209-212
//       void passall(bool), bool locked()      M17Demodulator.h:198-207
//
// Same names, same argument meaning.  Differences, all consequences of batching
// the per-sample loop onto the GPU:
//   * callbacks fire when a block of samples is drained (every `block` samples, on
//     flush() and in the destructor), not synchronously inside operator();
//   * the frame callback's bool result is ignored (the reference only uses it for the
//     unused DecodeResult of packet frames, M17FrameDecoder.h:306-311);
//   * samples must lie on the int16 grid `s / 41067.0` that apps/m17-demod.cpp:489
//     produces (the device consumes the int16 values);
//   * every instance behaves like a fresh m17-demod process (the reference's
//     function-local statics at M17Demodulator.h:422,659 are per instance here);
//   * there is no CPU fallback: construction throws std::runtime_error without a GPU.
#pragma once

#include "m17b.h"

#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

namespace mobilinkd
{

struct M17FrameDecoder
{
    enum class State { LSF, STREAM, BASIC_PACKET, FULL_PACKET, BERT };
    enum class SyncWordType { LSF, STREAM, PACKET, BERT };
    enum class DecodeResult { FAIL, OK, EOS, INCOMPLETE, PACKET_INCOMPLETE };
    enum class FrameType { LSF, LICH, STREAM, BASIC_PACKET, FULL_PACKET, BERT };

    using lsf_buffer_t = std::array<uint8_t, 30>;
    using lich_buffer_t = std::array<uint8_t, 6>;
    using audio_buffer_t = std::array<uint8_t, 18>;
    using packet_buffer_t = std::array<uint8_t, 26>;
    using bert_buffer_t = std::array<uint8_t, 25>;

    using output_buffer_t = struct {
        FrameType type;
        union {
            lich_buffer_t lich;
            audio_buffer_t stream;
            packet_buffer_t packet;
            bert_buffer_t bert;
        };
        lsf_buffer_t lsf;
    };

    using callback_t = std::function<bool(const output_buffer_t&, int)>;
};

template <typename FloatType>
struct M17Demodulator
{
    static constexpr uint16_t SAMPLE_RATE = 48000;
    static constexpr uint16_t SYMBOL_RATE = 4800;
    static constexpr uint16_t SAMPLES_PER_SYMBOL = SAMPLE_RATE / SYMBOL_RATE;
    static constexpr uint16_t BLOCK_SIZE = 192;

    using callback_t = M17FrameDecoder::callback_t;
    using diagnostic_callback_t = std::function<void(bool, FloatType, FloatType, FloatType, bool, FloatType, int, int, int, int)>;

    explicit M17Demodulator(callback_t callback, int device = 0, size_t block = 4800, uint32_t flags = 0)
    : callback_(std::move(callback)), block_(block)
    {
        m17b_config cfg{};
        cfg.n_streams = 1;
        cfg.device = device;
        cfg.max_chunk = uint32_t(block_);
        cfg.flags = flags | M17B_DIAG_RECORDS;      // the diagnostic callback fires at the reference cadence
        int rc = m17b_create(&cfg, &ctx_);
        if (rc != M17B_OK) throw std::runtime_error(std::string("m17b_create: ") + m17b_strerror(rc));
        pending_.reserve(block_);
    }

    M17Demodulator(const M17Demodulator&) = delete;
    M17Demodulator& operator=(const M17Demodulator&) = delete;

    virtual ~M17Demodulator()
    {
        try { flush(); } catch (...) {}
        m17b_destroy(ctx_);
    }

    bool locked() const { return dcd_; }
    void passall(bool enabled) { passall_ = enabled; }
    void diagnostics(diagnostic_callback_t callback) { diagnostic_callback = callback; }

    /// Reference signature: one scaled sample, `int16 / 41067.0`.
    void operator()(const FloatType input) { push(int16_t(std::lrint(double(input) * 41067.0))); }

    /// The same sample before scaling (what the device consumes).
    void push(int16_t sample)
    {
        pending_.push_back(sample);
        if (pending_.size() == block_) flush();
    }

    /// Demodulate everything buffered so far and fire the callbacks in the reference's order: a frame callback
    /// fires on the sample that completes the frame (M17Demodulator.h:646), the diagnostic callback every 384
    /// samples without carrier (:675-689) and every 960 with (:742-752), after the frame callback of the same sample.
    void flush()
    {
        if (!pending_.empty()) {
            int rc = m17b_submit(ctx_, pending_.data(), pending_.size(), pending_.size(), M17B_HOST);
            if (rc != M17B_OK) throw std::runtime_error(std::string("m17b_submit: ") + m17b_strerror(rc));
            submitted_ = true;      // the buffer must stay valid until the drain below; it is released right after it
        }
        frames_.clear();
        diags_.clear();
        size_t n = 0;
        int rc_drain = M17B_OK;
        do {
            const size_t at = frames_.size();
            frames_.resize(at + 256);
            rc_drain = m17b_drain(ctx_, frames_.data() + at, 256, &n);
            frames_.resize(at + (rc_drain == M17B_OK || rc_drain == M17B_EOVERFLOW ? n : 0));
        } while ((rc_drain == M17B_OK || rc_drain == M17B_EOVERFLOW) && n == 256);
        if (submitted_) { pending_.clear(); submitted_ = false; }     // consumed whatever the drain says: never resubmitted
        if (rc_drain != M17B_OK)
            throw std::runtime_error(std::string("m17b_drain: ") + m17b_strerror(rc_drain) + " " + m17b_last_cuda_error(ctx_));
        do {
            const size_t at = diags_.size();
            diags_.resize(at + 256);
            int rc = m17b_diag_drain(ctx_, diags_.data() + at, 256, &n);
            diags_.resize(at + (rc == M17B_OK ? n : 0));
            if (rc != M17B_OK) throw std::runtime_error(std::string("m17b_diag_drain: ") + m17b_strerror(rc));
        } while (n == 256);
        size_t fi = 0, di = 0;
        while (fi != frames_.size() || di != diags_.size()) {
            // the frame callback of sample p (0-based) precedes the diagnostic callback fired after p + 1 samples
            if (fi != frames_.size() && (di == diags_.size() || frames_[fi].sample_pos < diags_[di].d.samples)) {
                const m17b_frame& f = frames_[fi++];
                M17FrameDecoder::output_buffer_t out;
                out.type = M17FrameDecoder::FrameType(f.type);
                std::memcpy(out.packet.data(), f.payload, 26);
                std::memcpy(out.lsf.data(), f.lsf, 30);
                if (callback_) callback_(out, f.viterbi_cost);
            } else {
                const m17b_diag& d = diags_[di++].d;
                dcd_ = d.dcd;
                if (diagnostic_callback)
                    diagnostic_callback(d.dcd, FloatType(d.evm), FloatType(d.deviation), FloatType(d.offset), d.locked,
                                        FloatType(d.clock), d.sample_index, d.sync_index, d.clock_index, d.viterbi_cost);
            }
        }
    }

    diagnostic_callback_t diagnostic_callback;

private:
    callback_t callback_;
    size_t block_;
    m17b_ctx* ctx_ = nullptr;
    std::vector<int16_t> pending_;
    std::vector<m17b_frame> frames_;
    std::vector<m17b_diag_rec> diags_;
    bool submitted_ = false;
    bool dcd_ = false;
    bool passall_ = false;
};

} // mobilinkd
