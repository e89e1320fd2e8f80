// ref_demod.cpp -- C entry points around the UNMODIFIED reference headers.
//
// TEST INFRASTRUCTURE ONLY (see m17_oracle.h).  This translation unit contains
// no demodulator logic of its own: it #includes the reference's
// include/m17cxx/M17Demodulator.h from where it lies under /root/reference and
// drives M17Demodulator<float> exactly like apps/m17-demod.cpp:453-490 does
// (FloatType = float, `demod(sample / 41067.0)`), replacing only the codec2 /
// boost consumers with a callback that copies output_buffer_t into a record.
// Built by oracle/Makefile into oracle/_ref/ (git-ignored, never committed).
//
// The reference keeps two function-local statics (`initializing`,
// M17Demodulator.h:659, and `eot_flag`, :422), so ONE process can demodulate
// exactly ONE stream faithfully.  m17ref_demod_batch() therefore forks one
// child per stream -- the same "one m17-demod process per stream" model the
// CPU baseline uses.

#include "M17Demodulator.h"   // -I/root/reference/include/m17cxx -I/root/reference/blaze
#include "ax25_frame.h"       // the reference's AX.25 printer (used by apps/m17-demod.cpp:224-229)
#include <sstream>

#include "m17_oracle.h"

#include <cstdlib>
#include <cerrno>
#include <cstring>
#include <deque>
#include <new>
#include <cstdio>
#include <vector>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

bool display_lsf = false;     // required by M17FrameDecoder.h:19

using namespace mobilinkd;

namespace {

struct Sink {
    m17o_record* rec = nullptr;
    size_t cap = 0;
    size_t n = 0;
    uint32_t stream = 0;
    uint64_t pos = 0;
};

Sink* g_sink = nullptr;

bool on_frame(M17FrameDecoder::output_buffer_t const& f, int cost)
{
    Sink& s = *g_sink;
    if (s.n < s.cap) {
        m17o_record& r = s.rec[s.n];
        std::memset(&r, 0, sizeof(r));
        r.stream = s.stream;
        r.seq = uint32_t(s.n);
        r.sample_pos = s.pos;
        r.viterbi_cost = cost;
        r.type = uint8_t(f.type);
        static_assert(sizeof(f.packet) == 26);
        std::memcpy(r.payload, f.packet.data(), 26);   // union {lich,stream,packet,bert}
        std::memcpy(r.lsf, f.lsf.data(), 30);
    }
    s.n += 1;
    return true;
}

template <typename D>
void capture(D& d, m17o_trace& t)
{
    t.filtered = d.correlator.buffer_[d.correlator.prev_buffer_pos_];
    t.limit = d.correlator.limit();
    t.dcd_level = d.dcd.level();
    t.dev_offset = d.dev.offset();
    t.dev_idev = d.dev.idev();
    t.clock_est = d.clock_recovery.clock_estimate();
    t.sample_est = d.clock_recovery.sample_estimate_;
    t.evm = d.evm.evm();
    t.sync_count = d.sync_count;
    t.missing_sync_count = d.missing_sync_count;
    t.viterbi_cost = int(d.viterbi_cost);
    t.count = uint32_t(d.count_);
    t.demod_state = uint8_t(d.demodState);
    t.sample_index = d.sample_index;
    t.sync_sample_index = d.sync_sample_index;
    t.clock_sample_index = d.clock_recovery.sample_index();
    t.dcd = d.dcd_;
    t.dcd_triggered = d.dcd.dcd();
    t.sync_word_type = uint8_t(d.sync_word_type);
    t.decoder_state = uint8_t(d.decoder.state());
    t.corr_index = uint8_t(d.correlator.index());
    t.need_clock_reset = d.need_clock_reset_;
    t.need_clock_update = d.need_clock_update_;
    t.polarity = uint8_t(d.polarity);
}

} // namespace

extern "C" {

// Valid ONCE per process (reference statics).  apps/m17-demod.cpp:484-490.
size_t m17ref_demod_once(const int16_t* samples, size_t n, unsigned flags,
                         uint32_t stream_id, m17o_record* rec, size_t rec_cap,
                         m17o_trace* trace)
{
    Sink sink;
    sink.rec = rec; sink.cap = rec_cap; sink.stream = stream_id;
    g_sink = &sink;

    // The reference leaves several members indeterminate (M17FrameDecoder.h:102-104
    // output_buffer / depuncture_buffer, SyncWord::samples_); constructing the
    // unmodified class in zeroed storage makes the oracle deterministic.
    using Demod = M17Demodulator<float>;
    void* storage = std::calloc(1, sizeof(Demod));
    Demod& demod = *new (storage) Demod(on_frame);
    size_t total = n + ((flags & M17O_EOF_REPEAT) && n ? 1 : 0);
    for (size_t i = 0; i != total; ++i) {
        int16_t sample = samples[i < n ? i : n - 1];
        if (flags & M17O_INVERT) sample *= -1;
        sink.pos = i;
        demod(sample / 41067.0);
        if (trace) capture(demod, trace[i]);
    }
    g_sink = nullptr;
    demod.~Demod();
    std::free(storage);
    return sink.n;
}

// One forked child per stream, `jobs` children in flight.  Output arrays must be
// caller-allocated: rec[n_streams * rec_cap], counts[n_streams].
// trace (optional) = [n_streams * n_samples] entries.
int m17ref_demod_batch(const int16_t* samples, size_t n_streams, size_t n_samples,
                       size_t stride, unsigned flags, m17o_record* rec, size_t rec_cap,
                       uint64_t* counts, m17o_trace* trace, int jobs)
{
    if (jobs < 1) jobs = 1;
    size_t rec_bytes = n_streams * rec_cap * sizeof(m17o_record);
    size_t cnt_bytes = n_streams * sizeof(uint64_t);
    size_t trc_bytes = trace ? n_streams * n_samples * sizeof(m17o_trace) : 0;
    size_t total = rec_bytes + cnt_bytes + trc_bytes;
    if (total == 0) return 0;
    void* shm = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (shm == MAP_FAILED) return -1;
    auto* srec = reinterpret_cast<m17o_record*>(shm);
    auto* scnt = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(shm) + rec_bytes);
    auto* strc = trace ? reinterpret_cast<m17o_trace*>(reinterpret_cast<char*>(shm) + rec_bytes + cnt_bytes) : nullptr;

    // Children are reaped by pid, oldest first: a plain wait() would also collect unrelated children of
    // the calling process (bench.py's nvidia-smi sampler), miscount the live oracle processes and
    // return while one of them is still writing its records.
    std::deque<pid_t> pids;
    int status = 0, rc = 0;
    auto reap_oldest = [&]() {
        pid_t p = pids.front();
        pids.pop_front();
        pid_t r;
        do { r = waitpid(p, &status, 0); } while (r < 0 && errno == EINTR);
        if (r != p || !WIFEXITED(status) || WEXITSTATUS(status) != 0) rc = -2;
    };
    for (size_t s = 0; s != n_streams; ++s) {
        while ((int) pids.size() >= jobs) reap_oldest();
        pid_t pid = fork();
        if (pid < 0) { rc = -3; break; }
        if (pid == 0) {
            scnt[s] = m17ref_demod_once(samples + s * stride, n_samples, flags, uint32_t(s),
                                        srec + s * rec_cap, rec_cap,
                                        strc ? strc + s * n_samples : nullptr);
            _exit(0);
        }
        pids.push_back(pid);
    }
    while (!pids.empty()) reap_oldest();
    std::memcpy(rec, srec, rec_bytes);
    std::memcpy(counts, scnt, cnt_bytes);
    if (trace) std::memcpy(trace, strc, trc_bytes);
    munmap(shm, total);
    return rc;
}

// ---------------------------------------------------------------------------
// Component taps: each calls exactly one reference block so that the port (and
// the CUDA kernels) can be pinned block by block.  None of them has statics.
// ---------------------------------------------------------------------------

// CRC16.h:21-62.  reset(); for each byte operator()(b); return get().
uint16_t m17ref_crc16(const uint8_t* data, size_t n)
{
    CRC16<0x5935, 0xFFFF> crc;
    crc.reset();
    for (size_t i = 0; i != n; ++i) crc(data[i]);
    return crc.get();
}

uint32_t m17ref_golay_encode24(uint16_t data) { return Golay24::encode24(data); }

// Golay24.h:203-222
int m17ref_golay_decode(uint32_t input, uint32_t* output)
{
    uint32_t out = 0;
    bool ok = Golay24::decode(input, out);
    *output = out;
    return ok;
}

// Util.h:128-145 (llr<float,4>)
void m17ref_llr(float sample, int8_t* out2)
{
    auto r = llr<float, 4>(sample);
    out2[0] = std::get<0>(r);
    out2[1] = std::get<1>(r);
}

// Util.h:63-104: dump the 43-entry table (thresholds + LLR pairs).
void m17ref_llr_table(float* thresholds43, int8_t* pairs86)
{
    auto map = mobilinkd::detail::make_llr_map<float, 4>();
    for (size_t i = 0; i != map.size(); ++i) {
        thresholds43[i] = std::get<0>(map[i]);
        pairs86[2 * i] = std::get<0>(std::get<1>(map[i]));
        pairs86[2 * i + 1] = std::get<1>(std::get<1>(map[i]));
    }
}

// FirFilter.h:28-43 with the float RRC taps (M17Demodulator.h:79-118).
void m17ref_fir(const float* in, float* out, size_t n)
{
    BaseFirFilter<float, 150> f{mobilinkd::detail::Taps<float>::rrc_taps};
    for (size_t i = 0; i != n; ++i) out[i] = f(in[i]);
}

void m17ref_rrc_taps(float* taps150)
{
    for (size_t i = 0; i != 150; ++i) taps150[i] = mobilinkd::detail::Taps<float>::rrc_taps[i];
}

// apps/m17-demod.cpp:489 sample scaling as the reference does it.
float m17ref_scale(int16_t s) { return float(s / 41067.0); }

// DataCarrierDetect.h:53-73 as instantiated at M17Demodulator.h:149; update()
// every `period` samples.  level[i], dcd[i] after each update.
size_t m17ref_dcd(const float* in, size_t n, size_t period, float* level, uint8_t* trig)
{
    DataCarrierDetect<float, 48000, 400> dcd{2400, 3600, 0.1, 4.0};
    size_t k = 0;
    for (size_t i = 0; i != n; ++i) {
        dcd(in[i]);
        if ((i + 1) % period == 0) { dcd.update(); level[k] = dcd.level(); trig[k] = dcd.dcd(); ++k; }
    }
    return k;
}

// SlidingDFT.h:118-132 raw bins (re0, im0, re1, im1 per sample).
void m17ref_sdft(const float* in, size_t n, float* out4)
{
    NSlidingDFT<float, 48000, 120, 2> dft({2400, 3600});
    for (size_t i = 0; i != n; ++i) {
        auto r = dft(in[i]);
        out4[4 * i + 0] = r[0].real(); out4[4 * i + 1] = r[0].imag();
        out4[4 * i + 2] = r[1].real(); out4[4 * i + 3] = r[1].imag();
    }
}

// IirFilter.h:26-42 with the correlator's coefficients (Correlator.h:38-39).
void m17ref_limit_iir(const float* in, float* out, size_t n)
{
    Correlator<float> c;
    for (size_t i = 0; i != n; ++i) { c.sample(in[i]); out[i] = c.limit(); }
}

// Correlator.h:43-114: feed n samples, then report the 4 sync correlations and
// the outer symbol levels at `sample_index`.
void m17ref_correlator(const float* in, size_t n, uint8_t sample_index, float* corr4, float* mnmx2,
                       float* limit, uint8_t* index)
{
    Correlator<float> c;
    c.buffer_.fill(0);
    for (size_t i = 0; i != n; ++i) c.sample(in[i]);
    corr4[0] = c.correlate({+3,-3,+3,-3,+3,-3,+3,-3});
    corr4[1] = c.correlate({+3,+3,+3,+3,-3,-3,+3,-3});
    corr4[2] = c.correlate({3,-3,3,3,-3,-3,-3,-3});
    corr4[3] = c.correlate({+3,+3,+3,+3,+3,+3,-3,+3});
    auto [mn, mx] = c.outer_symbol_levels(sample_index);
    mnmx2[0] = mn; mnmx2[1] = mx;
    *limit = c.limit();
    *index = uint8_t(c.index());
}

// Correlator.h:43-64 per sample: out5[5 i + 0] = limit() after sample i, [1..4] = correlate() with the four
// sync words of M17Demodulator.h:154-157, correlator clocked on every sample.
void m17ref_correlate_trace(const float* in, size_t n, float* out5)
{
    Correlator<float> c;
    c.buffer_.fill(0);
    for (size_t i = 0; i != n; ++i) {
        c.sample(in[i]);
        out5[5 * i + 0] = c.limit();
        out5[5 * i + 1] = c.correlate({+3,-3,+3,-3,+3,-3,+3,-3});
        out5[5 * i + 2] = c.correlate({+3,+3,+3,+3,-3,-3,+3,-3});
        out5[5 * i + 3] = c.correlate({3,-3,3,3,-3,-3,-3,-3});
        out5[5 * i + 4] = c.correlate({+3,+3,+3,+3,+3,+3,-3,+3});
    }
}

// ax25_frame.h:22-263: what apps/m17-demod.cpp:224-229 prints for a reassembled packet.  Returns the text length.
size_t m17ref_ax25_print(const uint8_t* frame, size_t n, char* out, size_t cap)
{
    std::ostringstream os;
    write(os, ax25_frame(std::string(reinterpret_cast<const char*>(frame), n)));
    const std::string t = os.str();
    if (cap) { const size_t k = t.size() < cap - 1 ? t.size() : cap - 1; std::memcpy(out, t.data(), k); out[k] = 0; }
    return t.size();
}

// Util.h:320-413: PRBS9::validate over n bits (one per byte); `reset_before[i]` != 0 calls reset() first (what the
// CLI's diagnostic callback does when it sees no carrier).  out = {synced, bit_count, err_count}.
void m17ref_prbs9_validate(const uint8_t* bits, const uint8_t* reset_before, size_t n, uint32_t* out3)
{
    PRBS9 p;
    p.history.fill(0);
    for (size_t i = 0; i != n; ++i) {
        if (reset_before && reset_before[i]) p.reset();
        p.validate(bits[i] != 0);
    }
    out3[0] = p.synced; out3[1] = p.bit_count; out3[2] = p.err_count;
}

// KalmanFilter.h:18-66 + ClockRecovery.h:33-88.  ops[i]: 0 = reset(arg),
// 1 = update(uint8 arg) after `dt[i]` operator() calls, 2 = update() (dead
// reckoning) after dt[i] more calls.  out[i] = {sample_index, clock_estimate,
// sample_estimate}.
void m17ref_clock(const uint8_t* ops, const uint8_t* args, const uint32_t* dt, size_t n,
                  uint8_t* sample_index, float* clock_est, float* sample_est)
{
    ClockRecovery<float, 10> cr;
    for (size_t i = 0; i != n; ++i) {
        for (uint32_t k = 0; k != dt[i]; ++k) cr(0.f);
        switch (ops[i]) {
        case 0: cr.reset(args[i]); break;
        case 1: cr.update(args[i]); break;
        default: cr.update(); break;
        }
        sample_index[i] = cr.sample_index();
        clock_est[i] = cr.clock_estimate();
        sample_est[i] = cr.sample_estimate_;
    }
}

// FreqDevEstimator.h:26-53.  reset_before[i] != 0 -> dev.reset() first.
void m17ref_freqdev(const float* mn, const float* mx, const uint8_t* reset_before, size_t n,
                    float* offset, float* idev)
{
    FreqDevEstimator<float> dev;
    for (size_t i = 0; i != n; ++i) {
        if (reset_before[i]) dev.reset();
        dev.update(mn[i], mx[i]);
        offset[i] = dev.offset();
        idev[i] = dev.idev();
    }
}

// SymbolEvm.h:31-51
void m17ref_evm(const float* in, size_t n, float* out)
{
    SymbolEvm<float> evm;
    evm.reset();
    for (size_t i = 0; i != n; ++i) { evm.update(in[i]); out[i] = evm.evm(); }
}

// M17FrameDecoder.h:353-392 driven directly: frames[i] = 368 int8 soft bits,
// sync[i] = SyncWordType.  Records as for the demodulator.  cost_io carries
// viterbi_cost across frames (it is an in/out reference in the reference);
// costs[i]/states[i] report viterbi_cost and decoder.state() after frame i.
size_t m17ref_frame_decoder(const int8_t* frames, const uint8_t* sync, size_t n_frames,
                            m17o_record* rec, size_t rec_cap, int64_t* costs, uint8_t* states,
                            uint8_t* results)
{
    Sink sink;
    sink.rec = rec; sink.cap = rec_cap;
    g_sink = &sink;
    void* storage = std::calloc(1, sizeof(M17FrameDecoder));
    M17FrameDecoder& dec = *new (storage) M17FrameDecoder(on_frame);
    size_t cost = 0;
    for (size_t i = 0; i != n_frames; ++i) {
        M17FrameDecoder::input_buffer_t buf;
        std::memcpy(buf.data(), frames + i * 368, 368);
        sink.pos = i;
        auto r = dec(M17FrameDecoder::SyncWordType(sync[i]), buf, cost);
        costs[i] = int64_t(cost);
        states[i] = uint8_t(dec.state());
        results[i] = uint8_t(r);
    }
    g_sink = nullptr;
    dec.~M17FrameDecoder();
    std::free(storage);
    return sink.n;
}

// Viterbi.h:181-236 on an already depunctured buffer.  kind: 0 LSF 488->240,
// 1 stream 296->144, 2 packet 420->206, 3 bert 402->197.
size_t m17ref_viterbi(int kind, const int8_t* in, uint8_t* out)
{
    Trellis<4,2> trellis{makeTrellis<4, 2>({031,027})};
    Viterbi<decltype(trellis), 4> v{trellis};
    switch (kind) {
    case 0: { std::array<int8_t, 488> a; std::memcpy(a.data(), in, 488); std::array<uint8_t, 240> o; auto c = v.decode(a, o); std::memcpy(out, o.data(), 240); return c; }
    case 1: { std::array<int8_t, 296> a; std::memcpy(a.data(), in, 296); std::array<uint8_t, 144> o; auto c = v.decode(a, o); std::memcpy(out, o.data(), 144); return c; }
    case 2: { std::array<int8_t, 420> a; std::memcpy(a.data(), in, 420); std::array<uint8_t, 206> o; auto c = v.decode(a, o); std::memcpy(out, o.data(), 206); return c; }
    default: { std::array<int8_t, 402> a; std::memcpy(a.data(), in, 402); std::array<uint8_t, 197> o; auto c = v.decode(a, o); std::memcpy(out, o.data(), 197); return c; }
    }
}

// PolynomialInterleaver.h:21-24 / M17Randomizer.h:16-49 tables.
void m17ref_tables(uint16_t* interleave368, int8_t* dc368, int8_t* p1_61, int8_t* p2_12, int8_t* p3_8)
{
    PolynomialInterleaver<45, 92, 368> il;
    for (size_t i = 0; i != 368; ++i) interleave368[i] = uint16_t(il.index(i));
    M17Randomizer<368> r;
    for (size_t i = 0; i != 368; ++i) dc368[i] = r.dc_[i];
    for (size_t i = 0; i != 61; ++i) p1_61[i] = P1[i];
    for (size_t i = 0; i != 12; ++i) p2_12[i] = P2[i];
    for (size_t i = 0; i != 8; ++i) p3_8[i] = P3[i];
}

// SlidingDFT.h:85-95: the runtime std::exp() coefficients the DCD uses, read
// back through the public interface: on a fresh filter the first output is
// (0 + 1) * coeff == coeff exactly.
void m17ref_sdft_coeffs(float* re_im4)
{
    NSlidingDFT<float, 48000, 120, 2> dft({2400, 3600});
    auto r = dft(1.0f);
    re_im4[0] = r[0].real(); re_im4[1] = r[0].imag();
    re_im4[2] = r[1].real(); re_im4[3] = r[1].imag();
}

size_t m17ref_sizeof_demod() { return sizeof(M17Demodulator<float>); }

} // extern "C"
