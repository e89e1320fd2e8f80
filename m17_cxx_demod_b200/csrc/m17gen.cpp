// m17gen.cpp -- synthetic M17 baseband generator, CPU build (see include/m17gen.h; the logic is in
// m17gen_core.h, shared with the device-side modulator m17gen_dev.cu).  CPU only; not on the demodulation path.

#include "m17gen_core.h"

#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

using namespace m17gen;

extern "C" {

void m17gen_default_params(m17gen_params* p)
{
    std::memset(p, 0, sizeof(*p));
    p->seed = 0x4D3137;
    p->mode = M17GEN_STREAM;
    p->n_frames = 0;
    p->lead_in = 9600;
    p->tail = 4800;
    p->gap = 9600;
    p->snr_db = 40.f;
    p->gain = 1.f;
    p->phase = -1.f;
    p->idle_noise_rms = 8.f;
}

double m17gen_snr_linear(float snr_db) { return std::pow(10.0, double(snr_db) / 20.0); }

int m17gen_stream(const m17gen_params* p, int16_t* out, size_t n_samples, m17gen_truth* truth)
{
    if (!p || !out) return -1;
    if (truth) std::memset(truth, 0, sizeof(*truth));
    Rng rng(p->seed);
    const Plan q = make_plan(p, n_samples, rng);
    std::vector<int8_t> sym(q.n_sym);
    build_timeline(p, q, rng, sym.data(), truth);

    // the TX-domain samples once (the device build recomputes them per receiver sample: same values)
    std::vector<double> tx(q.tx_len + 8);
    double pw = 0; size_t pw_n = 0;
    for (size_t m = 0; m != tx.size(); ++m) {
        tx[m] = tx_sample(sym.data(), q.n_sym, q.amp, m);
        const size_t k_hi = m / 10;
        if (k_hi < q.n_sym && sym[k_hi]) { pw += tx[m] * tx[m]; ++pw_n; }      // == signal_rms()
    }
    const double sig_rms = pw_n ? std::sqrt(pw / double(pw_n)) : 0.0;
    const double sigma = noise_sigma(p, sig_rms, m17gen_snr_linear(p->snr_db));
    const double* txp = tx.data();
    const size_t txn = tx.size();
    for (size_t n = 0; n != n_samples; ++n)
        out[n] = rx_sample(p, q, sigma, n, [txp, txn](size_t m) { return m < txn ? txp[m] : 0.0; });
    return 0;
}

int m17gen_batch(const m17gen_params* params, size_t n_streams, int16_t* out, size_t n_samples,
                 size_t stride, m17gen_truth* truth, int n_threads)
{
    if (n_threads < 1) n_threads = 1;
    std::vector<std::thread> pool;
    std::vector<int> rcs(size_t(n_threads), 0);
    for (int t = 0; t != n_threads; ++t) {
        pool.emplace_back([&, t] {
            for (size_t s = size_t(t); s < n_streams; s += size_t(n_threads)) {
                int rc = m17gen_stream(&params[s], out + s * stride, n_samples, truth ? &truth[s] : nullptr);
                if (rc) rcs[size_t(t)] = rc;
            }
        });
    }
    for (auto& th : pool) th.join();
    for (int rc : rcs) if (rc) return rc;
    return 0;
}

uint16_t m17gen_crc16(const uint8_t* data, size_t n) { return crc16_m17(data, n); }
uint32_t m17gen_golay24_encode(uint16_t data12) { return golay24(data12); }
void m17gen_lsf_frame_bits(const uint8_t lsf[30], uint8_t out368[368]) { lsf_frame_bits(lsf, out368); }
void m17gen_stream_frame_bits(const uint8_t lsf[30], int lich_index, uint16_t fn, const uint8_t payload16[16],
                              uint8_t out368[368]) { stream_frame_bits(lsf, lich_index, fn, payload16, out368); }
void m17gen_packet_frame_bits(const uint8_t data26[26], uint8_t out368[368]) { packet_frame_bits(data26, out368); }
void m17gen_bert_frame_bits(uint16_t* prbs_state, uint8_t out368[368]) { bert_frame_bits(*prbs_state, out368); }
void m17gen_encode_callsign(const char* call, uint8_t out6[6]) { encode_callsign(call, out6); }

} // extern "C"
