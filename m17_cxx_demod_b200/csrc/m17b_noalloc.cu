// m17b_noalloc.cu -- m17b_filter_noalloc(): the matched filter (fir_kernel, kernel A) for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  m17b_debug_filter() runs fir_kernel
// into a temporary sample ring it cudaMallocs per call; this variant keeps that ring -- fir_kernel's scratch: ring_y,
// ring_x and the zero filter history -- in ONE file-scope __device__ array of a fixed, modest size, so the
// entry point allocates nothing: no cudaMalloc / cudaFree / cudaHostAlloc, no allocating library routine.
//
// The array's name is a host-side stand-in in host code, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress(), taken once per device with that device current (set_up()).  Every GPU that loads the library
// holds the array until exit.  All calls on a GPU share it and it keeps the last call's contents: what must start at
// zero (the filter history) is cleared on the caller's stream in every call; calls that overlap on one GPU race.
#include "m17b_dev.cuh"

#include <mutex>

using namespace m17b;

namespace {

constexpr uint32_t kMaxStreams = M17B_NOALLOC_MAX_STREAMS;     // = the padded stream count Np of the scratch ring
constexpr uint32_t kMaxSamples = M17B_NOALLOC_MAX_SAMPLES;
constexpr uint32_t kLines = 128;                                // ring lines: next_pow2((4000 + 31) / 32 + 1)
static_assert(kMaxStreams == 128 && (kMaxSamples + kLine - 1) / kLine + 1 <= kLines, "scratch ring size");

constexpr size_t kRingFloats = (size_t) kLines * kMaxStreams * kLine;
constexpr size_t kOffY = 0;
constexpr size_t kOffX = kOffY + kRingFloats * sizeof(float);
constexpr size_t kOffTail = kOffX + kRingFloats * sizeof(int16_t);
constexpr size_t kTailBytes = (size_t) kMaxStreams * kTailLen * sizeof(int16_t);
constexpr size_t kScratchBytes = kOffTail + kTailBytes;          // 3,186,688 bytes
static_assert(kScratchBytes == 3186688 && kOffX % 256 == 0 && kOffTail % 256 == 0, "256-byte aligned parts");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
struct PerDevice { bool ready = false; unsigned char* scratch = nullptr; };
PerDevice g_dev[kMaxDevices];
std::mutex g_mutex;
ConstTables g_tab;                  // host copies of the constant tables (static: nothing is allocated to build them)
uint32_t g_golay[2048];
bool g_tables_built = false;

// Once per device, with that device current: the scratch array's device address, the constant tables, kernel attributes.
int set_up(int device, unsigned char** scratch)
{
    std::lock_guard<std::mutex> lock(g_mutex);
    PerDevice& pd = g_dev[device];
    if (!pd.ready) {
        if (!g_tables_built) { build_tables(g_tab, g_golay); g_tables_built = true; }
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, g_noalloc_scratch) != cudaSuccess || !p) { cudaGetLastError(); return M17B_ECUDA; }
        if (upload_tables(g_tab, g_golay) != cudaSuccess) { cudaGetLastError(); return M17B_ECUDA; }
        if (configure_kernels() != cudaSuccess) { cudaGetLastError(); return M17B_ECUDA; }
        pd.scratch = static_cast<unsigned char*>(p);
        pd.ready = true;
    }
    *scratch = pd.scratch;
    return M17B_OK;
}

} // namespace

extern "C" int m17b_filter_noalloc(int device, const int16_t* samples_device, uint32_t n_streams, size_t n, size_t stream_stride,
                                   float* filtered_device, uint32_t flags, void* cuda_stream)
{
    if (!samples_device || !filtered_device || n_streams == 0 || n == 0 || stream_stride < n) return M17B_EINVAL;
    if (flags & ~(uint32_t) (M17B_FIR_FMA | M17B_INVERT)) return M17B_EINVAL;
    if (n_streams > kMaxStreams || n > kMaxSamples) return M17B_ETOOBIG;      // beyond the static scratch: use m17b_debug_filter
    if (device < 0 || device >= kMaxDevices || device >= m17b_device_count()) return M17B_ENODEV;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return M17B_ENODEV; }
    unsigned char* scratch = nullptr;
    int rc = set_up(device, &scratch);
    if (rc != M17B_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

    DevPtrs t{};
    t.n_streams = n_streams;
    t.np = kMaxStreams;
    uint32_t lines = 1;
    while (lines < (uint32_t) ((n + kLine - 1) / kLine + 1)) lines <<= 1;
    t.ring_lines = lines;                                            // <= kLines by the size check above
    t.ring_y = reinterpret_cast<float*>(scratch + kOffY);
    t.ring_x = reinterpret_cast<int16_t*>(scratch + kOffX);
    int16_t* tail = reinterpret_cast<int16_t*>(scratch + kOffTail);
    // the filter starts on zero history; the rings need no clearing (fir_kernel writes every position that is read back)
    if (cudaMemsetAsync(tail, 0, kTailBytes, st) != cudaSuccess) { cudaGetLastError(); return M17B_ECUDA; }
    launch_fir(t, samples_device, stream_stride, (uint32_t) n, tail, 0u, (flags & M17B_FIR_FMA) != 0, (flags & M17B_INVERT) != 0, st);
    launch_gather(t, (uint32_t) n, filtered_device, nullptr, st);
    if (cudaGetLastError() != cudaSuccess) return M17B_ECUDA;
    return M17B_OK;
}
