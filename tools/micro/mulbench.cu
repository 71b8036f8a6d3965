// Throughput of the 32x32 -> 64-bit multiply forms Philox needs, per SM sub-partition (B200).  Build: nvcc -arch=sm_100a -O3 mulbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(uint32_t* out, int iters, uint32_t seed) {
    uint32_t a[8];
    for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 8 + i;
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) { uint64_t p = (uint64_t)0xD2511F53u * a[i]; a[i] = (uint32_t)(p >> 32) ^ (uint32_t)p; }          // IMAD.WIDE.U32
            if (MODE == 1) { uint32_t h = __umulhi(0xD2511F53u, a[i]), l = 0xD2511F53u * a[i]; a[i] = h ^ l; }               // IMAD.HI.U32 + IMAD
            if (MODE == 2) { a[i] = 0xD2511F53u * a[i] + 12345u; }                                                              // IMAD (low 32)
            if (MODE == 3) { a[i] = __umulhi(0xD2511F53u, a[i]) + 1u; }                                                         // IMAD.HI.U32
            if (MODE == 4) { a[i] = (a[i] ^ seed) + (a[i] >> 3); }                                                              // ALU only
        }
    }
    for (int i = 0; i < 8; ++i) acc ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE> void run(const char* name) {
    uint32_t* d; cudaMalloc(&d, 148 * 8 * 1024 * 4);
    const int iters = 4096, blocks = 148 * 2, threads = 1024;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, threads>>>(d, 16, 1); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<MODE><<<blocks, threads>>>(d, iters, 1); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double warp_ops = (double)blocks * (threads / 32) * iters * 8;          // multiply "units" issued per warp
    double per_smsp_per_clk = warp_ops / (148.0 * 4) / (ms * 1e-3 * 1.965e9);
    printf("%-28s %.3f ms  %.3f warp-units/clk/SMSP  (= one unit every %.2f clk)\n", name, ms, per_smsp_per_clk, 1.0 / per_smsp_per_clk);
    cudaFree(d);
}
int main() {
    run<0>("IMAD.WIDE.U32 (+xor)"); run<1>("IMAD.HI.U32 + IMAD (+xor)"); run<2>("IMAD lo32"); run<3>("IMAD.HI.U32 (+add)"); run<4>("ALU xor/add/shift");
    return 0;
}
