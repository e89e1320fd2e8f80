// m17gen_dev.cu -- device-side synthetic M17 modulator (libm17gen_cuda.so; C ABI in include/m17gen.h).
//
// The same generator as csrc/m17gen.cpp (shared logic: m17gen_core.h), laid out for the GPU, so that every channel
// of a large run is its own waveform built where it is consumed (apps/m17-mod.cpp:204-224,264-308,
// 310-386,407-564 + this project's channel impairments):
//   timeline_kernel  one thread per stream: frame construction (LSF, LICH, stream / packet / BERT frames:
//                    convolutional code, puncturing, interleaver, randomiser, Golay, CRCs) -> symbol timeline,
//                    then the signal's RMS (a sequential sum, like the CPU's) -> noise sigma;
//   samples_kernel   one thread per output sample: RRC pulse shaping of the symbols around it, 6-point
//                    Lagrange resampling to the receiver clock, DC offset, AWGN (counter-based), quantisation.
// Integer work and IEEE double arithmetic only, compiled with -fmad=false: the int16 output is identical to the
// CPU generator's for the same parameters (tests/test_gpu_parity.py::test_device_modulator_matches_cpu_generator).
#include "m17gen_core.h"

#include <cuda_runtime.h>

using namespace m17gen;

namespace {

struct StreamPlan { Plan q; double sigma; };

__global__ void __launch_bounds__(64)
timeline_kernel(const m17gen_params* __restrict__ params, const double* __restrict__ snr_lin, uint32_t n_streams, size_t n_samples,
                int8_t* __restrict__ sym, size_t sym_stride, StreamPlan* __restrict__ plans)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    const m17gen_params* p = params + s;
    Rng rng(p->seed);
    const Plan q = make_plan(p, n_samples, rng);
    int8_t* my = sym + (size_t) s * sym_stride;
    build_timeline(p, q, rng, my, nullptr);
    plans[s].q = q;
    plans[s].sigma = noise_sigma(p, signal_rms(my, q), snr_lin[s]);
}

__global__ void __launch_bounds__(256)
samples_kernel(const m17gen_params* __restrict__ params, const StreamPlan* __restrict__ plans, const int8_t* __restrict__ sym,
               size_t sym_stride, size_t n_samples, int16_t* __restrict__ out, size_t stride)
{
    const uint32_t s = blockIdx.y;
    const m17gen_params* p = params + s;
    const StreamPlan pl = plans[s];
    const int8_t* my = sym + (size_t) s * sym_stride;
    for (size_t n = (size_t) blockIdx.x * blockDim.x + threadIdx.x; n < n_samples; n += (size_t) gridDim.x * blockDim.x) {
        const Plan& q = pl.q;
        out[(size_t) s * stride + n] = rx_sample(p, q, pl.sigma, n, [&](size_t m) { return tx_sample(my, q.n_sym, q.amp, m); });
    }
}

} // namespace

extern "C" int m17gen_batch_device(const m17gen_params* params, const double* snr_linear, size_t n_streams,
                                   int16_t* out_device, size_t n_samples, size_t stride, int device)
{
    if (!params || !snr_linear || !out_device || n_streams == 0 || n_samples == 0 || stride < n_samples) return -1;
    if (cudaSetDevice(device) != cudaSuccess) return -2;
    const size_t sym_stride = (max_symbols(n_samples) + 15) & ~size_t(15);
    m17gen_params* d_params = nullptr; double* d_snr = nullptr; int8_t* d_sym = nullptr; StreamPlan* d_plans = nullptr;
    int rc = 0;
    // streams in slices, so that the timelines of a very large batch need not all be resident at once
    const size_t slice = 16384;
    if (cudaMalloc(&d_params, slice * sizeof(m17gen_params)) != cudaSuccess || cudaMalloc(&d_snr, slice * sizeof(double)) != cudaSuccess ||
        cudaMalloc(&d_sym, slice * sym_stride) != cudaSuccess || cudaMalloc(&d_plans, slice * sizeof(StreamPlan)) != cudaSuccess) rc = -3;
    for (size_t s0 = 0; rc == 0 && s0 < n_streams; s0 += slice) {
        const size_t ns = n_streams - s0 < slice ? n_streams - s0 : slice;
        if (cudaMemcpy(d_params, params + s0, ns * sizeof(m17gen_params), cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemcpy(d_snr, snr_linear + s0, ns * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) { rc = -4; break; }
        timeline_kernel<<<(unsigned) ((ns + 63) / 64), 64>>>(d_params, d_snr, (uint32_t) ns, n_samples, d_sym, sym_stride, d_plans);
        const unsigned gx = (unsigned) ((n_samples + 255) / 256 < 64 ? (n_samples + 255) / 256 : 64);
        samples_kernel<<<dim3(gx, (unsigned) ns), 256>>>(d_params, d_plans, d_sym, sym_stride, n_samples, out_device + s0 * stride, stride);
        if (cudaDeviceSynchronize() != cudaSuccess) { rc = -5; break; }
    }
    cudaFree(d_params); cudaFree(d_snr); cudaFree(d_sym); cudaFree(d_plans);
    return rc;
}
