// Random-gather bandwidth probe: what DRAM/L2 can deliver for BVH-like access (random 80-B records
// out of a working set >> L2). Not part of the product; used to calibrate the roofline discussion.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t hash(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
template <int REC16, int STRIDE16>
__global__ void gather(const uint4 *__restrict__ buf, uint32_t nrec, int iters, uint32_t *out) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0, s = tid * 2654435761u + 12345u;
    for (int i = 0; i < iters; ++i) {
        s = hash(s + i);
        const uint4 *p = buf + (size_t)(s % nrec) * STRIDE16;
#pragma unroll
        for (int k = 0; k < REC16; ++k) { uint4 v = __ldg(p + k); acc += v.x ^ v.y ^ v.z ^ v.w; }
    }
    if (acc == 0x12345678u) out[0] = acc;
}
template <int REC16, int STRIDE16>
void run(const uint4 *buf, size_t bytes, uint32_t *out, const char *name) {
    uint32_t nrec = (uint32_t)(bytes / (16 * STRIDE16));
    int iters = 64;
    for (int bps : {8, 16}) {
        int grid = 148 * bps, block = 128;
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        gather<REC16, STRIDE16><<<grid, block>>>(buf, nrec, iters, out);
        cudaEventRecord(a);
        for (int r = 0; r < 5; ++r) gather<REC16, STRIDE16><<<grid, block>>>(buf, nrec, iters * 4, out);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        double recs = 5.0 * grid * block * iters * 4;
        printf("%-28s bps=%2d: %.1f Grec/s  useful %.0f GB/s  (ws %.0f MB)\n", name, bps, recs / ms / 1e6, recs * REC16 * 16 / ms / 1e6, bytes / 1e6);
    }
}
int main() {
    size_t bytes = 640ull << 20;
    uint4 *buf; uint32_t *out;
    cudaMalloc(&buf, bytes); cudaMalloc(&out, 4); cudaMemset(buf, 1, bytes);
    run<2, 2>(buf, bytes, out, "32B rec, 32B stride");
    run<4, 4>(buf, bytes, out, "64B rec, 64B stride");
    run<5, 5>(buf, bytes, out, "80B rec, 80B stride");
    run<5, 6>(buf, bytes, out, "80B rec, 96B stride");
    run<5, 8>(buf, bytes, out, "80B rec, 128B stride");
    run<8, 8>(buf, bytes, out, "128B rec, 128B stride");
    run<3, 3>(buf, bytes, out, "48B rec, 48B stride");
    run<5, 5>(buf, 64ull << 20, out, "80B rec, 80B stride L2-res");
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
