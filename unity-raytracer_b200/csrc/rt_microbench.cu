// rt_microbench.cu — measurement infrastructure, NOT part of the product library: two microbenchmarks that give bench.py
// the denominators of SURVEY.md §8(d)'s L2 and issue rooflines on the box it runs on (MEASURED_PEAKS.json holds only the
// HBM copy bandwidth and the bf16 tensor peak).  Built as librtb200_microbench.so next to librtb200.so.
//   rtmb_l2_read_gbs  : sustained read bandwidth over an L2-resident buffer (ld.global.cg: L1 bypassed, 16 B per lane)
//   rtmb_fp32_tflops  : sustained FP32 FMA rate (8 independent chains per thread)
#include <cuda_runtime.h>
#include <cstdint>

namespace {

__global__ void __launch_bounds__(256) k_l2_read(const uint4* __restrict__ buf, size_t n_vec, int reps, unsigned* sink) {
    unsigned acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
            uint4 v;
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    }
    if (acc == 0x9e3779b9u) *sink = acc;          // never true for the zero-filled buffer; keeps the loads alive
}

__global__ void __launch_bounds__(256) k_fma(float seed, int iters, float* sink) {
    float a0 = seed, a1 = seed + 1.f, a2 = seed + 2.f, a3 = seed + 3.f, a4 = seed + 4.f, a5 = seed + 5.f, a6 = seed + 6.f, a7 = seed + 7.f;
    const float m = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; ++i) {
        a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
        a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
    }
    float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123456.789f) *sink = s;
}

}  // namespace

extern "C" {

// returns GB/s (1e9 bytes per second), < 0 on a CUDA error
double rtmb_l2_read_gbs(int device, int mbytes, int reps) {
    if (cudaSetDevice(device) != cudaSuccess) return -1.0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1.0;
    const size_t bytes = (size_t)mbytes << 20, n_vec = bytes / 16;
    uint4* buf = nullptr; unsigned* sink = nullptr;
    if (cudaMalloc(&buf, bytes) != cudaSuccess || cudaMalloc(&sink, 4) != cudaSuccess) return -1.0;
    cudaMemset(buf, 0, bytes);
    const int grid = prop.multiProcessorCount * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_l2_read<<<grid, 256>>>(buf, n_vec, 2, sink);                         // warm the L2
    cudaEventRecord(e0);
    k_l2_read<<<grid, 256>>>(buf, n_vec, reps, sink);
    cudaEventRecord(e1);
    double out = -1.0;
    if (cudaEventSynchronize(e1) == cudaSuccess) {
        float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1);
        out = (double)bytes * reps / (ms * 1e-3) / 1e9;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf); cudaFree(sink);
    return out;
}

// returns TFLOP/s (an FMA = 2 flops), < 0 on a CUDA error
double rtmb_fp32_tflops(int device, int iters) {
    if (cudaSetDevice(device) != cudaSuccess) return -1.0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1.0;
    float* sink = nullptr;
    if (cudaMalloc(&sink, 4) != cudaSuccess) return -1.0;
    const int grid = prop.multiProcessorCount * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_fma<<<grid, 256>>>(1.0f, 1024, sink);
    cudaEventRecord(e0);
    k_fma<<<grid, 256>>>(1.0f, iters, sink);
    cudaEventRecord(e1);
    double out = -1.0;
    if (cudaEventSynchronize(e1) == cudaSuccess) {
        float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1);
        out = 2.0 * 8.0 * (double)iters * grid * 256 / (ms * 1e-3) / 1e12;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    return out;
}

}  // extern "C"
