// emu.cpp -- single-lane HOST build of the scan/decode lane logic.  TEST INFRASTRUCTURE ONLY.
//
// m17_cxx_demod_b200/csrc/m17b_lane.cuh is written without warp intrinsics, so the same source
// compiles here with g++ (-ffp-contract=off, baseline x86-64: the same separately rounded IEEE
// operations the kernels execute under -fmad=false).  This harness re-creates the device data
// layout in host memory, fills the sample ring with the matched filter in the reference's
// operation order, and then runs scan_lane()/decode_lane() for every stream in the same rounds
// m17b_submit()/m17b_sync() launch on the GPU.  tests/test_lane_emulation.py compares the frame
// records with the CPU oracle.  It lets the kernels' control flow (block/event passes, resume
// after a pending frame, ragged submits) be checked without a GPU; it is never shipped, and the
// product library does not contain it.
#include "m17b_lane.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace m17b {
ConstTables h_tab;
uint32_t h_golay[2048];
}

using namespace m17b;

namespace {

uint32_t next_pow2(uint32_t v) { uint32_t p = 1; while (p < v) p <<= 1; return p; }

struct Emu {
    DevPtrs d{};
    std::vector<uint32_t> st;
    std::vector<float> ring_y, corr_snap, sw_samples, fir_snap, seg_ring;
    std::vector<float2> ckpt_ring;
    std::vector<int16_t> llr, ring_x;
    std::vector<uint32_t> fix_list, dec_list, mask_ring;
    std::vector<uint8_t> dec_bytes, pkt_buf;
    std::vector<m17b_packet> pkt_recs;
    std::vector<m17b_frame> recs;
    Counters ctr{};
    LlrTable tab{};
    uint32_t head = 0;
    unsigned long long head64 = 0;
};

uint32_t g_origin = 0;       // ring position of the first sample (wrap-around tests)

void init(Emu& m, uint32_t n_streams, uint32_t max_chunk, size_t rec_cap, uint32_t ring_submits = 1)
{
    static bool tables = false;
    if (!tables) { build_tables(h_tab, h_golay); tables = true; }
    for (int i = 0; i != 48; ++i) m.tab.thr[i] = llr_thr_padded(h_tab.llr_thr, i);
    for (int i = 0; i != 43; ++i) {
        m.tab.val[i] = (uint32_t) (uint8_t) h_tab.llr_val[i][0] | ((uint32_t) (uint8_t) h_tab.llr_val[i][1] << 8);
    }
    DevPtrs& d = m.d;
    d.n_streams = n_streams;
    d.np = (n_streams + 127u) & ~127u;
    d.ring_lines = next_pow2((max_chunk * ring_submits + kRoundCap + 160 + 16 + kLine - 1) / kLine + 1);
    d.rec_cap = next_pow2((uint32_t) rec_cap);
    d.rec_read = 0;
    size_t np = d.np;
    m.st.assign((np / 32) * NFIELDS * 32, 0);
    m.ring_y.assign((size_t) d.ring_lines * np * kLine, 0.f);
    m.ring_x.assign((size_t) d.ring_lines * np * kLine, 0);
    m.corr_snap.assign(80 * np, 0.f);
    m.mask_ring.assign((size_t) d.ring_lines * np * 4, 0);
    m.ckpt_ring.assign((size_t) d.ring_lines * np * 4, float2{0.f, 0.f});
    d.n_seg = next_pow2((d.ring_lines * (uint32_t) kLine + 1920u) / (uint32_t) kSeg + 2u);
    m.seg_ring.assign((size_t) d.n_seg * 4 * np, 0.f);
    m.fix_list.assign(2 * np, 0);
    m.dec_list.assign(10 * np, 0);
    m.sw_samples.assign(30 * np, 0.f);
    m.fir_snap.assign((size_t) kTapsUsed * np, 0.f);
    m.llr.assign(184 * np, 0);
    m.dec_bytes.assign(56 * np, 0);
    m.recs.resize(d.rec_cap);
    m.pkt_buf.assign(np * (size_t) kPktStride, 0);
    d.pkt_cap = next_pow2(std::max<uint32_t>(64u, (uint32_t) np));
    m.pkt_recs.resize(d.pkt_cap);
    d.pkt_buf = m.pkt_buf.data(); d.pkt_recs = m.pkt_recs.data(); d.pkt_read = 0;
    d.st = m.st.data(); d.ring_y = m.ring_y.data(); d.ring_x = m.ring_x.data(); d.corr_snap = m.corr_snap.data();
    d.mask_ring = m.mask_ring.data(); d.ckpt_ring = m.ckpt_ring.data(); d.seg_ring = m.seg_ring.data(); d.fix_list = m.fix_list.data(); d.dec_list = m.dec_list.data(); d.sw_samples = m.sw_samples.data(); d.fir_snap = m.fir_snap.data();
    d.llr = m.llr.data(); d.dec_bytes = m.dec_bytes.data(); d.recs = m.recs.data(); d.ctr = &m.ctr;
    for (uint32_t s = 0; s != d.np; ++s) init_lane(d.st + st_index(s, 0), g_origin);
    m.head = g_origin;
}

void round(Emu& m, bool defer)
{
    DevPtrs& d = m.d;
    m.ctr.cls[1].lagging = 0;
    m.ctr.cls[1].deferred = 0;
    m.ctr.cls[1].fix_count = 0;
    for (int k = 0; k != 5; ++k) m.ctr.cls[1].dec_count[k] = 0;
    for (uint32_t s = 0; s != d.n_streams; ++s) scan_lane(d, make_io(d, s, &m.tab, 1), m.head, m.head64, true, defer);
    if (const char* t = getenv("M17EMU_TRACE")) {
        const uint32_t s = (uint32_t) atoi(t);
        const uint32_t* sb = d.st + st_index(s, 0);
        const uint32_t fl = sb[F_FLAGS * 32];
        fprintf(stderr, "emu s=%u pos=%u state=%u si=%u ssi=%u csi=%u dcd=%u trig=%u sc=%d miss=%d cnt=%u init=%u ckc=%u idx=%u fi=%u gk=%u evm=%.9g off=%.9g idev=%.9g ckest=%.9g h1=%.9g pend=%u fix=%u nr=%u nu=%u\n",
                s, sb[F_POS * 32], fl & 7, (fl >> FL_SI_SHIFT) & 15, (fl >> FL_SSI_SHIFT) & 15, (fl >> FL_CSI_SHIFT) & 15, !!(fl & FL_DCD), !!(fl & FL_DCD_TRIG),
                (int) sb[F_SYNC_COUNT * 32], (int) sb[F_MISSING * 32], sb[F_CNT * 32] & 0xFFFF, sb[F_CNT * 32] >> 16, sb[F_CK_COUNT * 32],
                sb[F_IDX * 32] & 0xFF, (sb[F_IDX * 32] >> 8) & 0xFFFF, sb[F_IDX * 32] >> 24, u2f(sb[F_EVM * 32]), u2f(sb[F_OFFSET * 32]), u2f(sb[F_IDEV * 32]),
                u2f(sb[F_CK_EST * 32]), u2f(sb[F_IIR0 * 32]), !!(fl & FL_PENDING), !!(fl & FL_FIXUP), !!(fl & FL_NEED_RESET), !!(fl & FL_NEED_UPDATE));
    }
    uint16_t hist[244];
    for (int k = 0; k != 5; ++k)
        for (uint32_t q = 0; q != m.ctr.cls[1].dec_count[k]; ++q) decode_lane(d, m.dec_list[(size_t) (5 + k) * d.np + q], hist, m.head, m.head64);
    for (uint32_t q = 0; q != m.ctr.cls[1].fix_count; ++q) {
        const uint32_t s = m.fix_list[d.np + q];
        fixup_lane(d, s, m.head);
        search_lane(d, s, m.head);
        d.st[st_index(s, F_FLAGS)] &= ~FL_FIXUP;
    }
}

// ---- M17B_STREAMING's two independent chains of rounds (m17b_api.cu: run_round with mask 1 / 2) -----------------
// classify_kernel's decision for one stream (m17b_kernels.cu; restated here because that kernel is warp code):
// the walk class the stream is listed in for a round of the chain `mask` (1 = receiving, 2 = acquiring), or -1.
int classify(const DevPtrs& d, uint32_t s, uint32_t head, uint32_t mask, bool defer, Counters& ctr)
{
    const uint32_t* sb = d.st + st_index(s, 0);
    const uint32_t fl = sb[F_FLAGS * 32], pos = sb[F_POS * 32];
    const int own = (mask & 1u) ? 0 : 1;
    if ((int32_t) (head - pos) <= 0) return -1;
    if (fl & (FL_PENDING | FL_FIXUP)) { ctr.cls[own].lagging += 1; return -1; }
    const int st = (int) ((fl >> FL_STATE_SHIFT) & 7);
    const bool receiving = (fl & FL_DCD) && (sb[F_CNT * 32] >> 16) == 0 && (sb[F_IDX * 32] >> 24) >= (uint32_t) kTapsUsed &&
                           (st == ST_FRAME || st == ST_STREAM_SYNC || st == ST_PACKET_SYNC || st == ST_BERT_SYNC || st == ST_SYNC_WAIT);
    const int cls = receiving ? 0 : 1;
    if (!((mask >> cls) & 1u)) return -1;
    if (defer && defers_at_once(sb, fl, pos, head)) { ctr.cls[own].deferred += 1; return -1; }
    return cls;
}

struct Task { int chain; uint32_t head; unsigned long long head64; uint32_t submit; };

// one round of one chain, as run_round(mask = 1 << chain) enqueues it
void chain_round(Emu& m, const Task& t, uint32_t guard)
{
    DevPtrs& d = m.d;
    const int c = t.chain;
    m.ctr.cls[c] = ClassCounters{};
    std::vector<uint32_t> list;
    for (uint32_t s = 0; s != d.n_streams; ++s)
        if (classify(d, s, t.head, 1u << c, true, m.ctr) == c) list.push_back(s);
    for (uint32_t s : list) scan_lane(d, make_io(d, s, &m.tab, c, c == 1), t.head, t.head64, true, true, guard);
    uint16_t hist[244];
    for (int k = 0; k != 5; ++k)
        for (uint32_t q = 0; q != m.ctr.cls[c].dec_count[k]; ++q)
            decode_lane(d, m.dec_list[(size_t) (c * 5 + k) * d.np + q], hist, t.head, t.head64);
    if (c == 1)
        for (uint32_t q = 0; q != m.ctr.cls[1].fix_count; ++q) {
            const uint32_t s = m.fix_list[d.np + q];
            fixup_lane(d, s, t.head);
            search_lane(d, s, t.head);
            d.st[st_index(s, F_FLAGS)] &= ~FL_FIXUP;
        }
}

} // namespace

extern "C" {

void m17emu_set_origin(uint32_t origin) { g_origin = origin; }

// The scan's soft-slicer look-up (llr4, m17b_lane.cuh) for n samples -> packed (second << 8 | first) pairs.
void m17emu_llr4(const float* samples, size_t n, uint16_t* out)
{
    Emu m;
    static bool tables = false;
    if (!tables) { build_tables(h_tab, h_golay); tables = true; }
    for (int i = 0; i != 48; ++i) m.tab.thr[i] = llr_thr_padded(h_tab.llr_thr, i);
    for (int i = 0; i != 43; ++i)
        m.tab.val[i] = (uint32_t) (uint8_t) h_tab.llr_val[i][0] | ((uint32_t) (uint8_t) h_tab.llr_val[i][1] << 8);
    for (size_t i = 0; i != n; ++i) out[i] = (uint16_t) llr4(m.tab, samples[i]);
}

// Demodulate [n_streams][n] int16 (row stride `stride`) fed in submits of the given sizes (cycled).
// Returns the number of records written to out (stream-major order is NOT guaranteed; seq is).
long m17emu_demod(const int16_t* samples, size_t n_streams, size_t n, size_t stride, const uint32_t* chunk_sizes,
                  size_t n_chunk_sizes, int invert, m17b_frame* out, size_t cap)
{
    const bool streaming = (invert & 2) != 0;
    const bool split = (invert & 4) != 0;          // streaming with the two chains of rounds run independently
    uint32_t lcg = (uint32_t) (invert >> 8) * 2654435761u + 12345u;      // interleaving seed
    invert &= 1;
    uint32_t max_chunk = 0;
    for (size_t i = 0; i != n_chunk_sizes; ++i) max_chunk = std::max(max_chunk, chunk_sizes[i]);
    Emu m;
    init(m, (uint32_t) n_streams, max_chunk, cap, split ? 4u : 1u);
    std::vector<Task> pending[2];                  // rounds enqueued and not yet run, per chain, in order
    uint32_t submit_no = 0;
    auto run_one = [&](int c) { chain_round(m, pending[c].front(), max_chunk); pending[c].erase(pending[c].begin()); };
    auto rnd = [&]() { lcg = lcg * 1664525u + 1013904223u; return lcg >> 16; };
    // matched filter over the whole capture in the reference's order (FirFilter.h:28-43)
    std::vector<float> x(n_streams * n), y(n_streams * n);
    for (size_t s = 0; s != n_streams; ++s) {
        float* xs = x.data() + s * n; float* ys = y.data() + s * n;
        for (size_t i = 0; i != n; ++i) {
            int v = samples[s * stride + i];
            if (invert) v = (int) (int16_t) (-v);
            xs[i] = (float) ((double) v / 41067.0);            // apps/m17-demod.cpp:489
        }
        for (size_t i = 0; i != n; ++i) {
            float r = 0.f;
            size_t kmax = std::min<size_t>(kTapsUsed - 1, i);
            for (size_t k = 0; k <= kmax; ++k) r = r + xs[i - k] * h_tab.taps[k];
            for (size_t k = kmax + 1; k < (size_t) kTapsUsed; ++k) r = r + 0.f * h_tab.taps[k];
            ys[i] = r;
        }
    }
    size_t off = 0, ci = 0;
    const DevPtrs& d = m.d;
    while (off < n) {
        uint32_t c = (uint32_t) std::min<size_t>(chunk_sizes[ci++ % n_chunk_sizes], n - off);
        if (split) {
            // the front end of submit k waits for the rounds of submit k - 2, both chains (m17b_submit: ev_rounds / ev_aux_rounds)
            for (int ch = 0; ch != 2; ++ch)
                while (!pending[ch].empty() && pending[ch].front().submit + 2 <= submit_no) run_one(ch);
        }
        for (size_t s = 0; s != n_streams; ++s)
            for (uint32_t i = 0; i != c; ++i) {
                size_t o = ring_off(m.head + i, d.ring_lines - 1, d.np, (uint32_t) s);
                size_t t = off + i;
                int v = samples[s * stride + t];
                if (invert) v = (int) (int16_t) (-v);
                m.ring_y[o] = y[s * n + t];
                m.ring_x[o] = (int16_t) v;
            }
        for (uint32_t s = 0; s != d.n_streams; ++s) {
            // dcd_lane's sample source (DcdSrc in m17b_kernels.cu), scalar: x[n] - x[n-120] from the int16 samples; before
            // the first sample of the stream the raw history is zero
            struct Src {
                const int16_t* row; long long off; uint32_t head_old, n_new; int invert; int rel0;
                float sx(long long t) const { int v = t >= 0 ? row[t] : 0; if (invert) v = (int) (int16_t) (-v); return scale_i16(v); }
                void request(uint32_t base) { rel0 = (int) (base - head_old); }
                void take(float (&dl)[8]) {
                    for (int j = 0; j != 8; ++j) {
                        const int r = rel0 + j;
                        dl[j] = (r >= 0 && (uint32_t) r < n_new) ? sx(off + r) - sx(off + r - kDftN) : 0.f;
                    }
                }
            } src{samples + s * stride, (long long) off, m.head, c, invert, 0};
            dcd_lane(d, s, 0, m.head, c, m.head64, src);
            dcd_lane(d, s, 1, m.head, c, m.head64, src);
        }
        m.head += c; m.head64 += c; off += c;
        int rounds = (int) (c / kMinFrameSpacing) + 2;
        if (split) {
            for (int ch = 0; ch != 2; ++ch)
                for (int r = 0; r != rounds; ++r) pending[ch].push_back(Task{ch, m.head, m.head64, submit_no});
            ++submit_no;
            // the two chains make progress in an arbitrary interleaving; either may fall behind by up to two submits
            for (uint32_t k = rnd() % (uint32_t) (4 * rounds + 1); k != 0; --k) {
                int ch = (int) (rnd() & 1u);
                if (pending[ch].empty()) ch ^= 1;
                if (pending[ch].empty()) break;
                run_one(ch);
            }
            continue;
        }
        // flags bit 1 = M17B_STREAMING: lanes may leave a partial frame for the next submit; only the
        // final flush (m17b_sync) brings every lane to the head
        for (int r = 0; r != rounds; ++r) round(m, streaming);
        for (int extra = 0; m.ctr.cls[1].lagging != 0; ++extra) {
            if (extra == 256) return -1;
            round(m, streaming);
        }
    }
    if (split) {
        while (!pending[0].empty() || !pending[1].empty()) {
            int ch = (int) (rnd() & 1u);
            if (pending[ch].empty()) ch ^= 1;
            run_one(ch);
        }
        round(m, false);                           // m17b_sync's flush rounds: one classification, every stream to the head
    }
    for (int extra = 0; m.ctr.cls[1].lagging != 0 || m.ctr.cls[1].deferred != 0; ++extra) {
        if (extra == 256) return -1;
        round(m, false);
    }
    if (m.ctr.dropped || m.ctr.faults) return -2;
    size_t k = std::min<size_t>(m.ctr.rec_count, cap);
    std::memcpy(out, m.recs.data(), k * sizeof(m17b_frame));
    return (long) k;
}

}
