// peaks.cu -- micro-benchmarks for the roofline denominators that MEASURED_PEAKS.json does not carry:
// register-resident FP64 and FP32 FMA throughput and L2-resident read bandwidth of the device the bench runs on.
// Measurement hooks only: nothing on the tracing path calls them.
//
//   the linear-scan kernels (k_closest_scan, k_shadow_scan, k_heavy_scan) are bound by the FP64 pipe: their roofline
//   fraction is DFMA-equivalent flop/s over the figure measured here;
//   the BVH kernels fetch nodes and leaf spheres through L1/L2: their tree bytes are quoted against the L2 figure.
#include <cuda_runtime.h>

#include <string>

#include "../../include/raytrace_b200.h"
#include "host_scene.h"

namespace {

constexpr int PEAK_THREADS = 256;
constexpr int PEAK_CHAINS = 8;  // independent accumulators per thread: enough to cover the FMA latency at 2+ warps per scheduler

template <typename T>
__global__ void __launch_bounds__(PEAK_THREADS) k_peak_fma(T* sink, int iters, T a, T b) {
    T x[PEAK_CHAINS];
#pragma unroll
    for (int k = 0; k < PEAK_CHAINS; k++) x[k] = (T)(threadIdx.x + k);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < PEAK_CHAINS; k++) x[k] = x[k] * a + b;  // contracted to one FMA (this file is built with -fmad=true)
    }
    T s = 0;
#pragma unroll
    for (int k = 0; k < PEAK_CHAINS; k++) s += x[k];
    if (s == (T)-1.2345) sink[0] = s;  // never true: keeps the chains alive
}

// every CTA walks the whole (L2-sized) buffer, 16 bytes per thread per step, through L2 only (ld.global.cg)
__global__ void __launch_bounds__(PEAK_THREADS) k_peak_l2(const uint4* buf, size_t n_vec, int rounds, uint32_t* sink) {
    uint32_t acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < rounds; r++)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
            const uint4 v = __ldcg(buf + i);
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345u) sink[0] = acc;
}

template <typename F>
cudaError_t time_ms(cudaStream_t st, F launch, float* ms) {
    cudaEvent_t e0, e1;
    cudaError_t e = cudaEventCreate(&e0);
    if (e != cudaSuccess) return e;
    e = cudaEventCreate(&e1);
    if (e != cudaSuccess) { cudaEventDestroy(e0); return e; }
    launch();  // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 3 && e == cudaSuccess; rep++) {
        cudaEventRecord(e0, st);
        launch();
        cudaEventRecord(e1, st);
        e = cudaEventSynchronize(e1);
        float t = 0;
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, e0, e1);
        if (t < best) best = t;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *ms = best;
    return e == cudaSuccess ? cudaGetLastError() : e;
}

}  // namespace

extern "C" int rt_measure_peaks(int device, rt_peaks* out) {
    using rt::set_error;
    if (!out) { set_error("rt_measure_peaks: null argument"); return RT_ERR_INVALID; }
    if (rt_device_count() <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_CUDA; }
    if (cudaSetDevice(device) != cudaSuccess) { set_error("cudaSetDevice failed"); return RT_ERR_CUDA; }
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    int l2_bytes = 0;
    cudaDeviceGetAttribute(&l2_bytes, cudaDevAttrL2CacheSize, device);
    cudaStream_t st;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { set_error("cudaStreamCreate failed"); return RT_ERR_CUDA; }
    void* sink = nullptr;
    uint4* buf = nullptr;
    // a quarter of the L2: resident whichever of the two dies' partitions a line is homed in
    const size_t l2_buf = ((size_t)(l2_bytes > 0 ? l2_bytes : (64 << 20)) / 4) & ~(size_t)4095;
    cudaError_t e = cudaMalloc(&sink, 64);
    if (e == cudaSuccess) e = cudaMalloc((void**)&buf, l2_buf);
    if (e == cudaSuccess) e = cudaMemsetAsync(buf, 1, l2_buf, st);
    float ms = 0;
    const int grid = sms * 8;  // 2048 threads per SM
    if (e == cudaSuccess) {
        const int iters = 1 << 13;
        e = time_ms(st, [&] { k_peak_fma<double><<<grid, PEAK_THREADS, 0, st>>>((double*)sink, iters, 1.0000001, 1e-9); }, &ms);
        out->fp64_tflops = 2.0 * PEAK_CHAINS * (double)iters * grid * PEAK_THREADS / (ms * 1e-3) / 1e12;
    }
    if (e == cudaSuccess) {
        const int iters = 1 << 15;
        e = time_ms(st, [&] { k_peak_fma<float><<<grid, PEAK_THREADS, 0, st>>>((float*)sink, iters, 1.0000001f, 1e-9f); }, &ms);
        out->fp32_tflops = 2.0 * PEAK_CHAINS * (double)iters * grid * PEAK_THREADS / (ms * 1e-3) / 1e12;
    }
    if (e == cudaSuccess) {
        const int rounds = 256;
        e = time_ms(st, [&] { k_peak_l2<<<grid, PEAK_THREADS, 0, st>>>(buf, l2_buf / 16, rounds, (uint32_t*)sink); }, &ms);
        out->l2_read_gbs = (double)l2_buf * rounds / (ms * 1e-3) / 1e9;
        out->l2_buffer_bytes = l2_buf;
    }
    out->sm_count = sms;
    cudaFree(sink); cudaFree(buf);
    cudaStreamDestroy(st);
    if (e != cudaSuccess) { set_error(std::string("rt_measure_peaks: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}
