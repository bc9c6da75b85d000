// l2peak.cu — measures the read bandwidth of the B200 L2 (MEASURED_PEAKS.json has HBM and tensor peaks only; SURVEY.md §8d asks
// for the L2 figure next to them). Every SM streams a buffer that fits the L2 (default 48 MiB of the 126 MB) with 128-bit loads,
// many times; the first pass (HBM -> L2) is not timed. Prints one JSON line. Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__global__ void __launch_bounds__(1024) stream_read(const uint4* __restrict__ buf, size_t n, int reps, unsigned* sink) {
    unsigned acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; r++)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            uint4 v;
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));   // .cg: L2 only
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345678u) *sink = acc;
}

int main(int argc, char** argv) {
    const size_t mib = argc > 1 ? (size_t)atoi(argv[1]) : 48;
    const size_t bytes = mib << 20, n = bytes / sizeof(uint4);
    uint4* buf; unsigned* sink;
    cudaMalloc(&buf, bytes); cudaMalloc(&sink, 4); cudaMemset(buf, 1, bytes);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int reps = 40, blocks = p.multiProcessorCount * 2;
    stream_read<<<blocks, 1024>>>(buf, n, 2, sink);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int t = 0; t < 5; t++) {
        cudaEventRecord(e0);
        stream_read<<<blocks, 1024>>>(buf, n, reps, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double gbs = (double)bytes * reps / (best * 1e-3) / 1e9;
    printf("{\"l2_read_gbs\": %.1f, \"buffer_mib\": %zu, \"l2_bytes\": %d, \"sms\": %d, \"reps\": %d, \"ms\": %.3f, \"error\": \"%s\"}\n", gbs, mib, p.l2CacheSize,
           p.multiProcessorCount, reps, best, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
