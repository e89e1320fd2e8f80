// m17b_api.cu -- host context behind the C ABI in include/m17b.h.
//
// One context = one GPU = N resident streams.  m17b_submit() enqueues
//     [H2D copy on the copy stream, double buffered]        (host input only)
//     fir_kernel over the new samples -> sample ring          (kernel A)
//     tail_kernel (keeps 160 raw samples of filter history)
//     dcd_kernel (carrier-detect pre-pass -> window sums)     (kernel A2)
//     R rounds of { scan_kernel ; decode_kernel ; fixup_kernel }   (kernels B, C)
// where a round lets every stream advance to its next completed frame (or by at
// most kRoundCap samples).  Frame completions of one stream are >= 1600 samples
// apart, so R = n/1600 + 2 rounds always reach the head; m17b_sync() verifies
// that on the device (`lagging`, `deferred`) and adds rounds if it ever is not so.
// Streams: `compute` (rounds; everything without M17B_STREAMING), `front` (filter,
// tail, pre-pass of the next submit under the rounds of the current one, with
// M17B_STREAMING), `copy` (sample upload), `d2h` (record read-back).
#include "m17b_dev.cuh"

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <new>
#include <cstdlib>
#include <string>
#include <vector>

using namespace m17b;

struct EvPair { cudaEvent_t a, b; int cat; };

struct m17b_ctx {
    m17b_config cfg{};
    DevPtrs d{};
    int16_t* tail[2] = {nullptr, nullptr};
    int cur_tail = 0;
    int16_t* stage[2] = {nullptr, nullptr};
    cudaEvent_t stage_free[2] = {nullptr, nullptr}, stage_ready[2] = {nullptr, nullptr};
    int cur_stage = 0;
    cudaStream_t compute = nullptr, copy = nullptr, d2h = nullptr;   // kernels; sample H2D; record D2H
    // M17B_STREAMING: the sample front end of submit k+1 (filter, tail, carrier-detect pre-pass) runs on `front`
    // while the rounds of submit k run on `compute`; ev_front = front end of the newest submit done,
    // ev_rounds[k & 1] = rounds of submit k done (the front end may be one submit ahead, not two)
    cudaStream_t front = nullptr;
    cudaStream_t aux = nullptr;                // chain of walk class 1 of a round (run_round)
    cudaEvent_t ev_cls = nullptr, ev_aux = nullptr;
    cudaEvent_t ev_aux_rounds[2] = {nullptr, nullptr};   // M17B_STREAMING: aux chain of submit k done (like ev_rounds)
    cudaEvent_t ev_front = nullptr, ev_rounds[2] = {nullptr, nullptr};
    unsigned long long n_submits = 0;
    uint32_t head = 0;
    unsigned long long head64 = 0;
    int sm_count = 0;
    std::string last_error;
    // frame records pulled from the device (pinned), waiting for m17b_drain
    m17b_frame* ready = nullptr;
    size_t ready_n = 0, ready_off = 0, ready_cap = 0;
    // profiling
    bool profile = false;
    std::vector<EvPair> ev_live, ev_pool;
    m17b_profile prof{};
    Counters host_ctr{};
    cudaEvent_t marks[8] = {};
    // one completion snapshot per submit in flight (m17b_drain_upto waits for one submit only)
    static constexpr int kSlots = 8;
    cudaEvent_t done[kSlots] = {};
    Counters* snap = nullptr;                  // pinned [kSlots]
    unsigned long long ticket = 0;             // submits issued
    unsigned long long slot_ticket[kSlots] = {};
    uint32_t rec_read = 0;                     // record cursor consumed by the host (monotonic, mod 2^32)
    unsigned long long dropped_reported = 0;   // overflow already signalled to the caller (M17B_EOVERFLOW is returned once per loss)
    uint32_t diag_read = 0;                    // diagnostic-record cursor consumed by the host
    uint32_t pkt_read = 0;                     // packet-record cursor consumed by the host
};

namespace {

enum { CAT_FIR, CAT_SCAN, CAT_DECODE, CAT_AUX, CAT_H2D, CAT_D2H, CAT_DCD, CAT_FIXUP, CAT_SCAN1 };

int fail(m17b_ctx* c, cudaError_t e, const char* what)
{
    if (c) c->last_error = std::string(what) + ": " + cudaGetErrorString(e);
    return e == cudaErrorMemoryAllocation ? M17B_ENOMEM : M17B_ECUDA;
}

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(c, e_, #call); } while (0)

uint32_t next_pow2(uint32_t v) { uint32_t p = 1; while (p < v) p <<= 1; return p; }

void ev_begin(m17b_ctx* c, cudaStream_t s, int cat, EvPair& p)
{
    if (!c->profile) return;
    if (!c->ev_pool.empty()) { p = c->ev_pool.back(); c->ev_pool.pop_back(); }
    else { cudaEventCreate(&p.a); cudaEventCreate(&p.b); }
    p.cat = cat;
    cudaEventRecord(p.a, s);
}

void ev_end(m17b_ctx* c, cudaStream_t s, EvPair& p)
{
    if (!c->profile) return;
    cudaEventRecord(p.b, s);
    c->ev_live.push_back(p);
}

void ev_collect(m17b_ctx* c)
{
    for (auto& p : c->ev_live) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
            switch (p.cat) {
            case CAT_FIR: c->prof.fir_ms += ms; break;
            case CAT_SCAN: c->prof.scan_ms += ms; break;
            case CAT_DECODE: c->prof.decode_ms += ms; break;
            case CAT_AUX: c->prof.aux_ms += ms; break;
            case CAT_H2D: c->prof.h2d_ms += ms; break;
            case CAT_DCD: c->prof.dcd_ms += ms; break;
            case CAT_FIXUP: c->prof.fixup_ms += ms; break;
            case CAT_SCAN1: c->prof.scan1_ms += ms; break;
            default: c->prof.d2h_ms += ms; break;
            }
        }
        c->ev_pool.push_back(p);
    }
    c->ev_live.clear();
}

// One round of one chain: the streams of the walk classes in `mask` that have samples to walk (classify_kernel)
// advance to their next completed frame (or until they need fixup_kernel), the completed frames are decoded, the
// fix-up / search requests served.  Class 0 = receiving streams, class 1 = acquiring ones (idle, start-up,
// sync-word search, preamble): few, but theirs are long serial walks.
//   mask 3 (m17b_sync's flush rounds, contexts without M17B_STREAMING): one classification on `st`, class 1's chain
//          forked to the aux stream beside class 0's, the round ends when both have;
//   mask 1 / 2 (M17B_STREAMING): the chain of one class alone on `st`.  The two chains of a submit then run
//          independently (m17b_submit), each classifying for itself: the flags word of a stream says whose it is,
//          and it is published last (scan_lane's write-back, decode_lane, fixup_done_kernel).
int run_round(m17b_ctx* c, bool defer, cudaStream_t st, uint32_t guard, uint32_t mask)
{
    EvPair p{};
    const int first = (mask & 1u) ? 0 : 1, last = (mask & 2u) ? 1 : 0;
    CK(cudaMemsetAsync(&c->d.ctr->cls[first], 0, (size_t) (last - first + 1) * sizeof(ClassCounters), st));
    launch_classify(c->d, c->head, mask, defer, st);
    const bool fork = mask == 3u;
    if (fork) {
        CK(cudaEventRecord(c->ev_cls, st));
        CK(cudaStreamWaitEvent(c->aux, c->ev_cls, 0));
    }
    for (int cls = last; cls >= first; --cls) {
        cudaStream_t s = (fork && cls) ? c->aux : st;
        ev_begin(c, s, cls ? CAT_SCAN1 : CAT_SCAN, p);
        launch_scan(c->d, c->head, c->head64, defer, guard, cls, mask == 2u, s);
        ev_end(c, s, p);
        ev_begin(c, s, CAT_DECODE, p);
        launch_decode(c->d, c->head, c->head64, cls, s);
        ev_end(c, s, p);
        c->prof.decode_launches += 1;
        if (cls) {                   // (a receiving stream that needs fixup_kernel becomes class 1 first, see scan_lane)
            ev_begin(c, s, CAT_FIXUP, p);
            launch_fixup(c->d, c->head, c->sm_count, cls, s);
            ev_end(c, s, p);
            c->prof.fixup_launches += 4;
        }
        c->prof.scan_launches += 1;
    }
    c->prof.scan_launches += 1;      // classify_kernel
    if (fork) {
        CK(cudaEventRecord(c->ev_aux, c->aux));
        CK(cudaStreamWaitEvent(st, c->ev_aux, 0));
    }
    CK(cudaGetLastError());
    return M17B_OK;
}

int fetch_counters(m17b_ctx* c)
{
    CK(cudaStreamSynchronize(c->aux));                       // the acquiring streams' chain (M17B_STREAMING) has its own stream
    CK(cudaMemcpyAsync(&c->host_ctr, c->d.ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->compute));
    CK(cudaStreamSynchronize(c->compute));
    return M17B_OK;
}

// Move the records [rec_read, cursor) off the device into c->ready (pinned staging) on the copy
// stream, so that the transfer overlaps kernels of later submits.  The device appends records at
// cursor % rec_cap with one monotonic atomic cursor, so records of one stream stay chronological.
int pull_records(m17b_ctx* c, uint32_t cursor)
{
    uint32_t n = cursor - c->rec_read;
    if (n == 0) return M17B_OK;
    if (n > c->d.rec_cap) n = c->d.rec_cap;                  // overflow: the device counted the rest as dropped
    if (c->ready_off == c->ready_n) c->ready_off = c->ready_n = 0;
    if (c->ready_n + n > c->ready_cap) {
        size_t cap = std::max<size_t>((c->ready_n + n) * 2, 4096);
        m17b_frame* nb = nullptr;
        CK(cudaHostAlloc((void**) &nb, cap * sizeof(m17b_frame), cudaHostAllocDefault));
        if (c->ready_n) std::memcpy(nb, c->ready, c->ready_n * sizeof(m17b_frame));
        if (c->ready) cudaFreeHost(c->ready);
        c->ready = nb;
        c->ready_cap = cap;
    }
    EvPair p{};
    // own stream: a record read-back never queues behind the next submit's sample upload
    ev_begin(c, c->d2h, CAT_D2H, p);
    const uint32_t first = c->rec_read & (c->d.rec_cap - 1);
    const uint32_t n1 = std::min(n, c->d.rec_cap - first);
    CK(cudaMemcpyAsync(c->ready + c->ready_n, c->d.recs + first, sizeof(m17b_frame) * n1, cudaMemcpyDeviceToHost, c->d2h));
    if (n1 != n)
        CK(cudaMemcpyAsync(c->ready + c->ready_n + n1, c->d.recs, sizeof(m17b_frame) * (n - n1), cudaMemcpyDeviceToHost, c->d2h));
    ev_end(c, c->d2h, p);
    CK(cudaStreamSynchronize(c->d2h));
    c->prof.d2h_bytes += sizeof(m17b_frame) * n;
    c->rec_read = cursor;
    c->d.rec_read = cursor;
    c->ready_n += n;
    return M17B_OK;
}

void free_all(m17b_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->cfg.device);
    if (c->compute) cudaStreamSynchronize(c->compute);
    if (c->copy) cudaStreamSynchronize(c->copy);
    if (c->d2h) cudaStreamSynchronize(c->d2h);
    if (c->front) cudaStreamSynchronize(c->front);
    if (c->aux) cudaStreamSynchronize(c->aux);
    for (auto& p : c->ev_live) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto& p : c->ev_pool) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto& e : c->marks) if (e) cudaEventDestroy(e);
    for (auto& e : c->done) if (e) cudaEventDestroy(e);
    if (c->snap) cudaFreeHost(c->snap);
    cudaFree(c->d.st); cudaFree(c->d.ring_y); cudaFree(c->d.ring_x); cudaFree(c->d.mask_ring); cudaFree(c->d.ckpt_ring);
    cudaFree(c->d.corr_snap); cudaFree(c->d.seg_ring); cudaFree(c->d.fix_list); cudaFree(c->d.scan_list); cudaFree(c->d.dec_list); cudaFree(c->d.diag_recs);
    cudaFree(c->d.sw_samples); cudaFree(c->d.fir_snap); cudaFree(c->d.llr); cudaFree(c->d.dec_bytes);
    cudaFree(c->d.recs); cudaFree(c->d.ctr); cudaFree(c->d.pkt_buf); cudaFree(c->d.pkt_recs);
    if (c->ready) cudaFreeHost(c->ready);
    for (int i = 0; i != 2; ++i) {
        cudaFree(c->tail[i]); cudaFree(c->stage[i]);
        if (c->stage_free[i]) cudaEventDestroy(c->stage_free[i]);
        if (c->stage_ready[i]) cudaEventDestroy(c->stage_ready[i]);
    }
    if (c->compute) cudaStreamDestroy(c->compute);
    if (c->copy) cudaStreamDestroy(c->copy);
    if (c->d2h) cudaStreamDestroy(c->d2h);
    if (c->front) cudaStreamDestroy(c->front);
    if (c->aux) cudaStreamDestroy(c->aux);
    if (c->ev_cls) cudaEventDestroy(c->ev_cls);
    if (c->ev_aux) cudaEventDestroy(c->ev_aux);
    for (auto& e : c->ev_aux_rounds) if (e) cudaEventDestroy(e);
    if (c->ev_front) cudaEventDestroy(c->ev_front);
    for (auto& e : c->ev_rounds) if (e) cudaEventDestroy(e);
    delete c;
}

} // namespace

extern "C" {

uint32_t m17b_version(void) { return M17B_VERSION; }

const char* m17b_strerror(int code)
{
    switch (code) {
    case M17B_OK: return "ok";
    case M17B_EINVAL: return "invalid argument";
    case M17B_ENODEV: return "no usable CUDA device";
    case M17B_ENOMEM: return "out of memory";
    case M17B_ECUDA: return "CUDA error";
    case M17B_EOVERFLOW: return "frame record ring overflow";
    case M17B_ESTATE: return "internal state error";
    case M17B_ETOOBIG: return "size exceeds the static scratch of m17b_filter_noalloc (at most 128 streams x 4000 samples per call)";
    default: return "unknown error";
    }
}

const char* m17b_last_cuda_error(const m17b_ctx* ctx) { return ctx ? ctx->last_error.c_str() : ""; }

int m17b_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int m17b_create(const m17b_config* cfg, m17b_ctx** out)
{
    if (!cfg || !out || cfg->n_streams == 0 || cfg->max_chunk == 0) return M17B_EINVAL;
    if (cfg->reserved[0] % 32u != 0 || cfg->reserved[1] || cfg->reserved[2]) return M17B_EINVAL;
    *out = nullptr;
    int ndev = m17b_device_count();
    if (ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) return M17B_ENODEV;
    m17b_ctx* c = new (std::nothrow) m17b_ctx;
    if (!c) return M17B_ENOMEM;
    c->cfg = *cfg;
    c->profile = cfg->flags & M17B_PROFILE;
    int rc = [&]() -> int {
        CK(cudaSetDevice(cfg->device));
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, cfg->device));
        c->sm_count = prop.multiProcessorCount;
        {
            // the rounds are a chain of latency-bound launches: highest priority, so that their CTAs are placed
            // as soon as CTAs of a concurrent filter launch (front stream, M17B_STREAMING) retire
            int prio_lo = 0, prio_hi = 0;
            CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
            CK(cudaStreamCreateWithPriority(&c->compute, cudaStreamNonBlocking, prio_hi));
            CK(cudaStreamCreateWithPriority(&c->front, cudaStreamNonBlocking, prio_lo));
            CK(cudaStreamCreateWithPriority(&c->aux, cudaStreamNonBlocking, prio_hi));
            CK(cudaEventCreateWithFlags(&c->ev_cls, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&c->ev_aux, cudaEventDisableTiming));
            for (auto& e : c->ev_aux_rounds) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&c->ev_front, cudaEventDisableTiming));
            for (auto& e : c->ev_rounds) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        }
        CK(cudaStreamCreateWithFlags(&c->copy, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&c->d2h, cudaStreamNonBlocking));

        DevPtrs& d = c->d;
        d.n_streams = cfg->n_streams;
        d.np = (cfg->n_streams + 127u) & ~127u;
        // ring must hold one submit plus the lag a stream may keep plus the look-back
        // of the slow filter path / sliding DFT
        // M17B_STREAMING: the next submit's filter writes its chunk while the lanes of this one are still
        // running their rounds (front stream), so a lane may be a whole chunk further behind the newest ring write
        // (... and an acquiring stream is served by a chain of its own, at its own pace, and changes chains a few times
        // on its way into a transmission: it may trail the head by up to two more submits, hence four in all)
        uint32_t need = cfg->max_chunk * ((cfg->flags & M17B_STREAMING) ? 4u : 1u) + kRoundCap + 160 + 16;
        d.ring_lines = next_pow2((need + kLine - 1) / kLine + 1);
        // at most 2 callbacks (LICH + LSF) per completed frame, completions >= kMinFrameSpacing apart; room for
        // three submits: the one being drained and two enqueued behind it (m17b_drain_upto pipelines of depth <= 2;
        // deeper pipelines pass their own frame_capacity)
        d.rec_cap = cfg->frame_capacity ? cfg->frame_capacity
                                        : (uint32_t) std::min<uint64_t>((uint64_t) cfg->n_streams * 6 * (cfg->max_chunk / kMinFrameSpacing + 2), 0x7fffffffull);
        d.rec_cap = next_pow2(std::max(d.rec_cap, 64u));          // slot = cursor % rec_cap must survive cursor wrap-around
        d.rec_read = 0;
        for (auto& e : c->done) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        CK(cudaHostAlloc((void**) &c->snap, sizeof(Counters) * m17b_ctx::kSlots, cudaHostAllocDefault));
        std::memset(c->snap, 0, sizeof(Counters) * m17b_ctx::kSlots);
        size_t np = d.np;
        size_t ring_bytes = (size_t) d.ring_lines * np * kLine * sizeof(float);
        CK(cudaMalloc(&d.st, (np / 32) * NFIELDS * 32 * sizeof(uint32_t)));
        CK(cudaMalloc(&d.ring_y, ring_bytes));
        CK(cudaMalloc(&d.ring_x, ring_bytes / 2));
        CK(cudaMalloc(&d.mask_ring, (size_t) d.ring_lines * np * 4 * sizeof(uint32_t)));
        CK(cudaMalloc(&d.ckpt_ring, (size_t) d.ring_lines * np * 4 * sizeof(float2)));
        CK(cudaMalloc(&d.corr_snap, 80 * np * sizeof(float)));
        if (cfg->flags & M17B_DIAG_RECORDS) {
            // one record per 384 samples and stream at most; room for three submits like the frame records
            d.diag_cap = next_pow2((uint32_t) std::min<uint64_t>((uint64_t) cfg->n_streams * 3 * (cfg->max_chunk / 384 + 2), 0x40000000ull));
            CK(cudaMalloc(&d.diag_recs, (size_t) d.diag_cap * sizeof(m17b_diag_rec)));
        }
        d.pkt_cap = next_pow2(std::max(64u, cfg->n_streams));
        CK(cudaMalloc(&d.pkt_buf, np * (size_t) kPktStride));
        CK(cudaMalloc(&d.pkt_recs, (size_t) d.pkt_cap * sizeof(m17b_packet)));
        // a scan lane may lag the head by the ring span; its window starts up to 960 samples earlier, and the
        // first window's entry must survive 1920 samples longer
        d.n_seg = next_pow2((d.ring_lines * (uint32_t) kLine + 1920u) / (uint32_t) kSeg + 2u);
        CK(cudaMalloc(&d.seg_ring, (size_t) d.n_seg * 4 * np * sizeof(float)));
        CK(cudaMemsetAsync(d.seg_ring, 0, (size_t) d.n_seg * 4 * np * sizeof(float), c->compute));
        CK(cudaMalloc(&d.fix_list, kClasses * np * sizeof(uint32_t)));
        CK(cudaMalloc(&d.scan_list, kClasses * np * sizeof(uint32_t)));
        CK(cudaMalloc(&d.dec_list, kClasses * 5 * np * sizeof(uint32_t)));
        CK(cudaMalloc(&d.sw_samples, 30 * np * sizeof(float)));
        CK(cudaMalloc(&d.fir_snap, (size_t) kTapsUsed * np * sizeof(float)));
        CK(cudaMalloc(&d.llr, 184 * np * sizeof(int16_t)));
        CK(cudaMalloc(&d.dec_bytes, 56 * np));
        CK(cudaMalloc(&d.recs, (size_t) d.rec_cap * sizeof(m17b_frame)));
        CK(cudaMalloc(&d.ctr, sizeof(Counters)));
        CK(cudaMemsetAsync(d.ring_y, 0, ring_bytes, c->compute));
        CK(cudaMemsetAsync(d.ring_x, 0, ring_bytes / 2, c->compute));
        CK(cudaMemsetAsync(d.mask_ring, 0, (size_t) d.ring_lines * np * 4 * sizeof(uint32_t), c->compute));
        CK(cudaMemsetAsync(d.ckpt_ring, 0, (size_t) d.ring_lines * np * 4 * sizeof(float2), c->compute));
        CK(cudaMemsetAsync(d.corr_snap, 0, 80 * np * sizeof(float), c->compute));
        CK(cudaMemsetAsync(d.sw_samples, 0, 30 * np * sizeof(float), c->compute));
        CK(cudaMemsetAsync(d.fir_snap, 0, (size_t) kTapsUsed * np * sizeof(float), c->compute));
        CK(cudaMemsetAsync(d.llr, 0, 184 * np * sizeof(int16_t), c->compute));
        CK(cudaMemsetAsync(d.dec_bytes, 0, 56 * np, c->compute));
        CK(cudaMemsetAsync(d.ctr, 0, sizeof(Counters), c->compute));
        for (int i = 0; i != 2; ++i) {
            CK(cudaMalloc(&c->tail[i], (size_t) cfg->n_streams * kTailLen * sizeof(int16_t)));
            CK(cudaMemsetAsync(c->tail[i], 0, (size_t) cfg->n_streams * kTailLen * sizeof(int16_t), c->compute));
            CK(cudaEventCreateWithFlags(&c->stage_free[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&c->stage_ready[i], cudaEventDisableTiming));
        }
        ConstTables tab;
        std::vector<uint32_t> golay(2048);
        build_tables(tab, golay.data());
        CK(upload_tables(tab, golay.data()));
        CK(configure_kernels());
        // debug hook (tests): the stream's first sample sits at ring position `origin` instead of 0, so that a
        // short run crosses the 2^32 wrap of the 32-bit ring positions
        c->head = cfg->reserved[0];
        launch_init_state(d, c->head, c->compute);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(c->compute));
        return M17B_OK;
    }();
    if (rc != M17B_OK) {
        std::fprintf(stderr, "m17b_create: %s\n", c->last_error.c_str());
        free_all(c);
        return rc;
    }
    *out = c;
    return M17B_OK;
}

void m17b_destroy(m17b_ctx* ctx) { free_all(ctx); }

int m17b_host_alloc(void** ptr, size_t bytes)
{
    if (!ptr) return M17B_EINVAL;
    cudaError_t e = cudaHostAlloc(ptr, bytes, cudaHostAllocDefault);
    return e == cudaSuccess ? M17B_OK : M17B_ENOMEM;
}

void m17b_host_free(void* ptr) { if (ptr) cudaFreeHost(ptr); }

int m17b_device_alloc(m17b_ctx* c, void** ptr, size_t bytes)
{
    if (!c || !ptr) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMalloc(ptr, bytes));
    return M17B_OK;
}

void m17b_device_free(m17b_ctx* c, void* ptr)
{
    if (!c || !ptr) return;
    cudaSetDevice(c->cfg.device);
    cudaStreamSynchronize(c->compute);
    cudaFree(ptr);
}

int m17b_memcpy_h2d(m17b_ctx* c, void* dst_device, const void* src_host, size_t bytes)
{
    if (!c) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpy(dst_device, src_host, bytes, cudaMemcpyHostToDevice));
    return M17B_OK;
}

int m17b_submit(m17b_ctx* c, const int16_t* samples, size_t n, size_t stride, int location)
{
    if (!c || !samples || n == 0 || n > c->cfg.max_chunk || stride < n) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    const uint32_t ns = c->cfg.n_streams;
    const int16_t* dev_in = samples;
    size_t dev_stride = stride;
    EvPair p{};
    // M17B_STREAMING: the sample front end (filter, tail, carrier-detect pre-pass) has its own stream and runs
    // while the rounds of the previous submit are still going: it touches no scan state, and with the ring
    // doubled it writes no ring position a lane of the previous submit still reads (ring guard in scan_lane).
    // It may be one submit ahead of the rounds, not two.
    const bool streaming = (c->cfg.flags & M17B_STREAMING) != 0;
    cudaStream_t fs = streaming ? c->front : c->compute;
    if (streaming) {                                                                    // rounds of submit k-2, both chains
        CK(cudaStreamWaitEvent(fs, c->ev_rounds[c->n_submits & 1], 0));
        CK(cudaStreamWaitEvent(fs, c->ev_aux_rounds[c->n_submits & 1], 0));
    }
    if (location == M17B_HOST) {
        int b = c->cur_stage;
        c->cur_stage ^= 1;
        if (!c->stage[b]) CK(cudaMalloc(&c->stage[b], (size_t) ns * c->cfg.max_chunk * sizeof(int16_t)));
        CK(cudaStreamWaitEvent(c->copy, c->stage_free[b], 0));       // kernels of two submits ago are done with it
        ev_begin(c, c->copy, CAT_H2D, p);
        if (stride == n)      // contiguous [N][n]: one linear copy (full PCIe rate, no per-row descriptors)
            CK(cudaMemcpyAsync(c->stage[b], samples, (size_t) ns * n * sizeof(int16_t), cudaMemcpyHostToDevice, c->copy));
        else
            CK(cudaMemcpy2DAsync(c->stage[b], n * sizeof(int16_t), samples, stride * sizeof(int16_t),
                                 n * sizeof(int16_t), ns, cudaMemcpyHostToDevice, c->copy));
        ev_end(c, c->copy, p);
        CK(cudaEventRecord(c->stage_ready[b], c->copy));
        CK(cudaStreamWaitEvent(fs, c->stage_ready[b], 0));
        c->prof.h2d_bytes += (uint64_t) ns * n * sizeof(int16_t);
        dev_in = c->stage[b];
        dev_stride = n;
    }
    {
        const bool invert = c->cfg.flags & M17B_INVERT;
        ev_begin(c, fs, CAT_FIR, p);
        launch_fir(c->d, dev_in, dev_stride, (uint32_t) n, c->tail[c->cur_tail], c->head, c->cfg.flags & M17B_FIR_FMA, invert, fs);
        ev_end(c, fs, p);
        ev_begin(c, fs, CAT_AUX, p);
        launch_tail(c->d, dev_in, dev_stride, (uint32_t) n, c->tail[c->cur_tail], c->tail[c->cur_tail ^ 1], fs);
        ev_end(c, fs, p);
    }
    // carrier-detect pre-pass over the new samples (reads the same int16 input and the raw history the filter read)
    ev_begin(c, fs, CAT_DCD, p);
    launch_dcd(c->d, dev_in, dev_stride, c->tail[c->cur_tail], c->head, (uint32_t) n, c->head64, c->cfg.flags & M17B_INVERT, fs);
    ev_end(c, fs, p);
    if (location == M17B_HOST) CK(cudaEventRecord(c->stage_free[c->cur_stage ^ 1], fs));      // the staging buffer may be refilled
    c->prof.dcd_launches += 1;
    if (streaming) {
        CK(cudaEventRecord(c->ev_front, fs));
        CK(cudaStreamWaitEvent(c->compute, c->ev_front, 0));
        CK(cudaStreamWaitEvent(c->aux, c->ev_front, 0));
    }
    CK(cudaGetLastError());
    c->cur_tail ^= 1;
    c->prof.fir_launches += 1;
    c->prof.aux_launches += 1;
    {
        uint32_t b0 = c->head & ~7u;
        uint32_t span = c->head + (uint32_t) n - b0;
        c->prof.fir_samples += (uint64_t) c->d.np * (((span + 127) / 128) * 128);
    }
    c->head += (uint32_t) n;
    c->head64 += n;
    c->prof.scan_samples += (uint64_t) ns * n;

    // Rounds.  A stream completes at most one frame per round.  M17B_STREAMING: streams walk whole frames
    // (scan_lane's deferral), so ceil(n / 1920) rounds serve every stream in steady reception; the remaining
    // rounds let stragglers (acquisition, carrier-detect gaps, fix-ups) catch up.  A round costs a frame's
    // worth of serial latency however few lanes run; the next submit's front end fills the machine meanwhile.
    const int rounds = (int) (n / kMinFrameSpacing) + 2;
    if (streaming) {
        // two independent chains: the receiving streams' rounds on `compute`, the acquiring streams' on `aux`.  An
        // acquiring stream's round is a long serial walk (a preamble, a search span, then a frame) and there are few
        // such streams, so they get about half as many rounds, run beside the others; they may trail the head by
        // more than a submit meanwhile (the ring is doubled; `faults` guards it).
        const uint32_t guard = c->cfg.max_chunk;
        for (int r = 0; r != rounds; ++r) {
            int rc = run_round(c, true, c->compute, guard, 1u);
            if (rc != M17B_OK) return rc;
        }
        for (int r = 0; r != rounds; ++r) {
            int rc = run_round(c, true, c->aux, guard, 2u);
            if (rc != M17B_OK) return rc;
        }
        CK(cudaEventRecord(c->ev_aux_rounds[c->n_submits & 1], c->aux));
    } else {
        for (int r = 0; r != rounds; ++r) {
            int rc = run_round(c, false, c->compute, 0u, 3u);
            if (rc != M17B_OK) return rc;
        }
    }
    if (streaming) CK(cudaEventRecord(c->ev_rounds[c->n_submits & 1], c->compute));
    c->n_submits += 1;
    // completion snapshot of this submit: counters (incl. the record cursor) + an event
    const int slot = (int) (c->ticket % m17b_ctx::kSlots);
    CK(cudaMemcpyAsync(&c->snap[slot], c->d.ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->compute));
    CK(cudaEventRecord(c->done[slot], c->compute));
    c->ticket += 1;
    c->slot_ticket[slot] = c->ticket;
    return M17B_OK;
}

int m17b_sync(m17b_ctx* c)
{
    if (!c) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    int rc = fetch_counters(c);
    if (rc != M17B_OK) return rc;
    auto behind = [&] { return c->host_ctr.cls[0].lagging | c->host_ctr.cls[0].deferred | c->host_ctr.cls[1].lagging | c->host_ctr.cls[1].deferred; };
    for (int extra = 0; behind() != 0; ++extra) {
        if (extra == 256) { c->last_error = "streams failed to reach the head"; return M17B_ESTATE; }
        if ((rc = run_round(c, false, c->compute, 0u, 3u)) != M17B_OK) return rc;
        if ((rc = fetch_counters(c)) != M17B_OK) return rc;
    }
    ev_collect(c);
    if (c->host_ctr.dropped != c->dropped_reported) {       // signalled once per loss; the surviving records stay drainable
        c->dropped_reported = c->host_ctr.dropped;
        return M17B_EOVERFLOW;
    }
    return M17B_OK;
}

int m17b_pending_frames(m17b_ctx* c, size_t* n)
{
    if (!c || !n) return M17B_EINVAL;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    *n = (c->ready_n - c->ready_off) + std::min(c->host_ctr.rec_count - c->rec_read, c->d.rec_cap);
    return rc;
}

int m17b_drain(m17b_ctx* c, m17b_frame* out, size_t cap, size_t* n)
{
    if (!c || !n || (!out && cap)) return M17B_EINVAL;
    *n = 0;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    int rc2 = pull_records(c, c->host_ctr.rec_count);
    if (rc2 != M17B_OK) return rc2;
    size_t k = std::min(cap, c->ready_n - c->ready_off);
    if (k) std::memcpy(out, c->ready + c->ready_off, k * sizeof(m17b_frame));
    c->ready_off += k;
    *n = k;
    c->prof.decode_frames += k;
    return rc;
}

unsigned long long m17b_ticket(const m17b_ctx* c) { return c ? c->ticket : 0; }

int m17b_drain_upto(m17b_ctx* c, unsigned long long ticket, m17b_frame* out, size_t cap, size_t* n)
{
    if (!c || !n || (!out && cap)) return M17B_EINVAL;
    *n = 0;
    bool overflow = false;
    CK(cudaSetDevice(c->cfg.device));
    if (c->ready_off == c->ready_n && ticket != 0) {
        if (ticket > c->ticket) return M17B_EINVAL;
        // the newest snapshot slot that is not newer than `ticket` and has not been recycled
        unsigned long long t = ticket;
        if (c->ticket - t >= (unsigned long long) m17b_ctx::kSlots) t = c->ticket - m17b_ctx::kSlots + 1;
        const int slot = (int) ((t - 1) % m17b_ctx::kSlots);
        CK(cudaEventSynchronize(c->done[slot]));
        const Counters& k = c->snap[slot];
        int rc = pull_records(c, k.rec_count);
        if (rc != M17B_OK) return rc;
        if (k.dropped != c->dropped_reported) { c->dropped_reported = k.dropped; overflow = true; }
    }
    size_t k = std::min(cap, c->ready_n - c->ready_off);
    if (k) std::memcpy(out, c->ready + c->ready_off, k * sizeof(m17b_frame));
    c->ready_off += k;
    *n = k;
    c->prof.decode_frames += k;
    return overflow ? M17B_EOVERFLOW : M17B_OK;
}

int m17b_diag_drain(m17b_ctx* c, m17b_diag_rec* out, size_t cap, size_t* n)
{
    if (!c || !n || (!out && cap)) return M17B_EINVAL;
    *n = 0;
    if (!c->d.diag_recs) return M17B_EINVAL;                // context created without M17B_DIAG_RECORDS
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    uint32_t avail = c->host_ctr.diag_count - c->diag_read;
    if (avail > c->d.diag_cap) { c->diag_read = c->host_ctr.diag_count - c->d.diag_cap; avail = c->d.diag_cap; }
    const uint32_t k = (uint32_t) std::min<size_t>(cap, avail);
    const uint32_t first = c->diag_read & (c->d.diag_cap - 1);
    const uint32_t n1 = std::min(k, c->d.diag_cap - first);
    if (n1) CK(cudaMemcpy(out, c->d.diag_recs + first, sizeof(m17b_diag_rec) * n1, cudaMemcpyDeviceToHost));
    if (k != n1) CK(cudaMemcpy(out + n1, c->d.diag_recs, sizeof(m17b_diag_rec) * (k - n1), cudaMemcpyDeviceToHost));
    c->diag_read += k;
    c->d.diag_read = c->diag_read;
    *n = k;
    return c->host_ctr.diag_dropped ? M17B_EOVERFLOW : M17B_OK;
}

int m17b_diag_get(m17b_ctx* c, uint32_t stream, m17b_diag* out)
{
    if (!c || !out || stream >= c->cfg.n_streams) return M17B_EINVAL;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    uint32_t f[NFIELDS];
    CK(cudaMemcpy2D(f, sizeof(uint32_t), c->d.st + st_index(stream, 0), 32 * sizeof(uint32_t), sizeof(uint32_t),
                    NFIELDS, cudaMemcpyDeviceToHost));
    auto asf = [&](int i) { float v; std::memcpy(&v, &f[i], 4); return v; };
    uint32_t fl = f[F_FLAGS];
    // arguments of diagnostic_callback at M17Demodulator.h:746-750
    out->dcd = (fl & FL_DCD) ? 1 : 0;
    // the evm argument is SymbolEvm::evm() (StandardDeviation.h:79-82) at the update points reached with the carrier
    // detect on (:746) and dev.error() == 0 (FreqDevEstimator.h:53) at those reached with it off (:683)
    out->evm = (fl & FL_DCD) ? std::sqrt(asf(F_EVM)) : 0.f;
    out->deviation = 2400.f / asf(F_IDEV);                                // FreqDevEstimator.h:52
    out->offset = asf(F_OFFSET);
    out->locked = ((fl >> FL_STATE_SHIFT) & 7) != ST_UNLOCKED;
    out->clock = asf(F_CK_EST);
    out->sample_index = (int) ((fl >> FL_SI_SHIFT) & 15);
    out->sync_index = (int) ((fl >> FL_SSI_SHIFT) & 15);
    out->clock_index = (int) ((fl >> FL_CSI_SHIFT) & 15);
    out->viterbi_cost = (int) f[F_VCOST];
    out->samples = c->head64 - (uint64_t) (c->head - f[F_POS]);
    return M17B_OK;
}

int m17b_counters_get(m17b_ctx* c, m17b_counters* out)
{
    if (!c || !out) return M17B_EINVAL;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    out->samples = c->head64 * c->cfg.n_streams;
    for (int i = 0; i != 6; ++i) out->frames[i] = c->host_ctr.frames[i];
    out->lsf_crc_fail = c->host_ctr.lsf_crc_fail;
    out->golay_fail = c->host_ctr.golay_fail;
    out->faults = c->host_ctr.faults;
    out->dropped = c->host_ctr.dropped;
    out->packets_ok = c->host_ctr.packets_ok;
    out->packets_crc_fail = c->host_ctr.packets_crc_fail;
    out->packet_seq_errors = c->host_ctr.packet_seq_errors;
    out->packets_dropped = c->host_ctr.packets_dropped;
    return M17B_OK;
}

int m17b_packet_drain(m17b_ctx* c, m17b_packet* out, size_t cap, size_t* n)
{
    if (!c || !n || (!out && cap)) return M17B_EINVAL;
    *n = 0;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    uint32_t avail = c->host_ctr.pkt_count - c->pkt_read;
    if (avail > c->d.pkt_cap) { c->pkt_read = c->host_ctr.pkt_count - c->d.pkt_cap; avail = c->d.pkt_cap; }
    const uint32_t k = (uint32_t) std::min<size_t>(cap, avail);
    const uint32_t first = c->pkt_read & (c->d.pkt_cap - 1);
    const uint32_t n1 = std::min(k, c->d.pkt_cap - first);
    if (n1) CK(cudaMemcpy(out, c->d.pkt_recs + first, sizeof(m17b_packet) * n1, cudaMemcpyDeviceToHost));
    if (k != n1) CK(cudaMemcpy(out + n1, c->d.pkt_recs, sizeof(m17b_packet) * (k - n1), cudaMemcpyDeviceToHost));
    c->pkt_read += k;
    c->d.pkt_read = c->pkt_read;
    *n = k;
    return M17B_OK;
}

int m17b_ber_get(m17b_ctx* c, uint32_t stream, m17b_ber* out)
{
    if (!c || !out || stream >= c->cfg.n_streams) return M17B_EINVAL;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    uint32_t f[NFIELDS];
    CK(cudaMemcpy2D(f, sizeof(uint32_t), c->d.st + st_index(stream, 0), 32 * sizeof(uint32_t), sizeof(uint32_t),
                    NFIELDS, cudaMemcpyDeviceToHost));
    out->synced = (int32_t) ((f[F_PRBS] >> 9) & 1u);
    out->bits = f[F_BER_BITS];
    out->errors = f[F_BER_ERRS];
    out->packet_bytes = f[F_PKT] & 0xFFFFu;
    return M17B_OK;
}

int m17b_get_profile(m17b_ctx* c, m17b_profile* out, int reset)
{
    if (!c || !out) return M17B_EINVAL;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    *out = c->prof;
    out->search_requests = c->host_ctr.search_requests - c->prof.search_requests;    // prof.* hold the values at the last reset
    out->search_samples = c->host_ctr.search_samples - c->prof.search_samples;
    if (reset) {
        c->prof = m17b_profile{};
        c->prof.search_requests = c->host_ctr.search_requests;
        c->prof.search_samples = c->host_ctr.search_samples;
    }
    return M17B_OK;
}

int m17b_profile_enable(m17b_ctx* c, int on)
{
    if (!c) return M17B_EINVAL;
    int rc = m17b_sync(c);
    if (rc != M17B_OK && rc != M17B_EOVERFLOW) return rc;
    c->profile = on != 0;
    return M17B_OK;
}

int m17b_event_record(m17b_ctx* c, int slot)
{
    if (!c || slot < 0 || slot >= 8) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    if (!c->marks[slot]) CK(cudaEventCreate(&c->marks[slot]));
    CK(cudaEventRecord(c->marks[slot], c->compute));
    return M17B_OK;
}

int m17b_event_elapsed(m17b_ctx* c, int a, int b, double* ms)
{
    if (!c || !ms || a < 0 || a >= 8 || b < 0 || b >= 8 || !c->marks[a] || !c->marks[b]) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaEventSynchronize(c->marks[b]));
    float f = 0;
    CK(cudaEventElapsedTime(&f, c->marks[a], c->marks[b]));
    *ms = f;
    return M17B_OK;
}

static int debug_filter(m17b_ctx* c, const int16_t* samples_host, size_t n, size_t stride, float* filtered_host,
                        float* scaled_host, float* corr_host)
{
    if (!c || !samples_host || !filtered_host || n == 0 || stride < n) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    const uint32_t ns = c->cfg.n_streams;
    DevPtrs t{};
    t.n_streams = ns;
    t.np = c->d.np;
    t.ring_lines = next_pow2((uint32_t) ((n + kLine - 1) / kLine + 1));
    size_t ring_bytes = (size_t) t.ring_lines * t.np * kLine * sizeof(float);
    int16_t *din = nullptr, *dtail = nullptr;
    float *dy = nullptr, *dx = nullptr, *dcorr = nullptr;
    int rc = [&]() -> int {
        CK(cudaMalloc(&t.ring_y, ring_bytes));
        CK(cudaMalloc(&t.ring_x, ring_bytes / 2));
        CK(cudaMalloc(&din, (size_t) ns * n * sizeof(int16_t)));
        CK(cudaMalloc(&dtail, (size_t) ns * kTailLen * sizeof(int16_t)));
        CK(cudaMalloc(&dy, (size_t) ns * n * sizeof(float)));
        CK(cudaMalloc(&dx, (size_t) ns * n * sizeof(float)));
        CK(cudaMemsetAsync(dtail, 0, (size_t) ns * kTailLen * sizeof(int16_t), c->compute));
        CK(cudaMemcpy2DAsync(din, n * sizeof(int16_t), samples_host, stride * sizeof(int16_t), n * sizeof(int16_t), ns,
                             cudaMemcpyHostToDevice, c->compute));
        launch_fir(t, din, n, (uint32_t) n, dtail, 0, c->cfg.flags & M17B_FIR_FMA, c->cfg.flags & M17B_INVERT, c->compute);
        if (corr_host) {
            CK(cudaMalloc(&dcorr, (size_t) ns * n * 5 * sizeof(float)));
            launch_debug_corr(t, (uint32_t) n, dcorr, c->compute);
            CK(cudaMemcpyAsync(corr_host, dcorr, (size_t) ns * n * 5 * sizeof(float), cudaMemcpyDeviceToHost, c->compute));
        }
        launch_gather(t, (uint32_t) n, dy, dx, c->compute);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(filtered_host, dy, (size_t) ns * n * sizeof(float), cudaMemcpyDeviceToHost, c->compute));
        if (scaled_host)
            CK(cudaMemcpyAsync(scaled_host, dx, (size_t) ns * n * sizeof(float), cudaMemcpyDeviceToHost, c->compute));
        CK(cudaStreamSynchronize(c->compute));
        return M17B_OK;
    }();
    cudaFree(t.ring_y); cudaFree(t.ring_x); cudaFree(din); cudaFree(dtail); cudaFree(dy); cudaFree(dx); cudaFree(dcorr);
    return rc;
}

int m17b_debug_filter(m17b_ctx* c, const int16_t* samples_host, size_t n, size_t stride, float* filtered_host,
                      float* scaled_host)
{
    return debug_filter(c, samples_host, n, stride, filtered_host, scaled_host, nullptr);
}

int m17b_debug_correlate(m17b_ctx* c, const int16_t* samples_host, size_t n, size_t stride, float* filtered_host,
                         float* corr_host)
{
    if (!corr_host) return M17B_EINVAL;
    return debug_filter(c, samples_host, n, stride, filtered_host, nullptr, corr_host);
}

int m17b_measure_fp32_peak(m17b_ctx* c, int variant, double* tflops)
{
    if (!c || !tflops) return M17B_EINVAL;
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->compute));
    double v = run_fp32_peak(variant, c->sm_count, c->compute);
    if (v < 0) return fail(c, cudaGetLastError(), "fp32 peak kernel");
    *tflops = v;
    return M17B_OK;
}

} // extern "C"
