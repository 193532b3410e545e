// peaks.cu — measures the roofline denominators SURVEY.md 8(d) asks for and MEASURED_PEAKS.json does not hold:
// L2-resident streaming read bandwidth, HBM streaming read bandwidth (cross-check of hbm_gbs) and the FP32 FMA rate.
// Stand-alone tool (lib/gsb_peaks), run on the GPU box:  gsb_peaks > gpurun_out/peaks.json
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } \
    } while (0)

// every thread streams 128-bit loads over the buffer `reps` times; the xor keeps the loads alive
__global__ void __launch_bounds__(256) k_read(const uint4* __restrict__ buf, size_t n16, int reps, uint32_t* sink) {
    uint4 acc = make_uint4(0, 0, 0, 0);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; r++) {
        size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < n16; i += 4 * stride) {
            const uint4 a = __ldcg(buf + i), b = __ldcg(buf + i + stride), c = __ldcg(buf + i + 2 * stride), d = __ldcg(buf + i + 3 * stride);
            acc.x ^= a.x ^ b.x ^ c.x ^ d.x; acc.y ^= a.y ^ b.y ^ c.y ^ d.y;
            acc.z ^= a.z ^ b.z ^ c.z ^ d.z; acc.w ^= a.w ^ b.w ^ c.w ^ d.w;
        }
        for (; i < n16; i += stride) {
            const uint4 a = __ldcg(buf + i);
            acc.x ^= a.x; acc.y ^= a.y; acc.z ^= a.z; acc.w ^= a.w;
        }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x12345678u) *sink = acc.x;
}

// 8 independent FMA chains per thread
__global__ void __launch_bounds__(256) k_fma(int iters, float seed, float* sink) {
    float a[8];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = seed + (float)(threadIdx.x + k);
    const float m = 1.0000001f, c = 1e-7f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = fmaf(a[k], m, c);
    }
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; k++) s += a[k];
    if (s == 12345.678f) *sink = s;
}

static float timeIt(cudaStream_t s, int repeats, const auto& launch) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < repeats; r++) {
        CK(cudaEventRecord(e0, s));
        launch();
        CK(cudaEventRecord(e1, s));
        CK(cudaEventSynchronize(e1));
        float ms = 0.0f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        best = std::min(best, ms);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return best;
}

int main() {
    int dev = 0, sms = 0, clockKHz = 0, l2Bytes = 0;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaDeviceGetAttribute(&clockKHz, cudaDevAttrClockRate, dev));
    CK(cudaDeviceGetAttribute(&l2Bytes, cudaDevAttrL2CacheSize, dev));
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    uint32_t* sink = nullptr;
    CK(cudaMalloc(&sink, 64));
    const int grid = sms * 8;

    // L2-resident: 32 MiB (well inside one die's share of the 126 MB L2), warmed, read 64 times per launch
    const size_t l2Buf = 32ull << 20;
    uint4* small = nullptr;
    CK(cudaMalloc(&small, l2Buf));
    CK(cudaMemset(small, 1, l2Buf));
    const int l2Reps = 64;
    k_read<<<grid, 256, 0, s>>>(small, l2Buf / 16, 2, sink);
    const float l2Ms = timeIt(s, 10, [&] { k_read<<<grid, 256, 0, s>>>(small, l2Buf / 16, l2Reps, sink); });
    const double l2Gbs = (double)l2Buf * l2Reps / (l2Ms * 1e-3) / 1e9;

    // HBM: 8 GiB, read once per launch
    const size_t hbmBuf = 8ull << 30;
    uint4* big = nullptr;
    CK(cudaMalloc(&big, hbmBuf));
    CK(cudaMemset(big, 1, hbmBuf));
    k_read<<<grid, 256, 0, s>>>(big, hbmBuf / 16, 1, sink);
    const float hbmMs = timeIt(s, 6, [&] { k_read<<<grid, 256, 0, s>>>(big, hbmBuf / 16, 1, sink); });
    const double hbmGbs = (double)hbmBuf / (hbmMs * 1e-3) / 1e9;

    // FP32 FMA
    const int iters = 1 << 16;
    float* fs = reinterpret_cast<float*>(sink) + 4;
    k_fma<<<grid, 256, 0, s>>>(1024, 1.0f, fs);
    const float fmaMs = timeIt(s, 6, [&] { k_fma<<<grid, 256, 0, s>>>(iters, 1.0f, fs); });
    const double fmaTflops = (double)grid * 256 * 8 * 2 * (double)iters / (fmaMs * 1e-3) / 1e12;

    printf("{\"sms\": %d, \"sm_clock_mhz\": %.0f, \"l2_bytes\": %d, \"l2_read_gbs\": %.1f, \"l2_buffer_mib\": %zu, "
           "\"hbm_read_gbs\": %.1f, \"hbm_buffer_gib\": %zu, \"fp32_fma_tflops\": %.2f, "
           "\"fp32_theoretical_tflops\": %.2f, \"how\": \"128-bit ld.global.cg streams, best of N, CUDA events; 8 FMA chains per thread\"}\n",
           sms, clockKHz / 1e3, l2Bytes, l2Gbs, l2Buf >> 20, hbmGbs, hbmBuf >> 30, fmaTflops, sms * 128.0 * 2.0 * clockKHz * 1e3 / 1e12);
    return 0;
}
