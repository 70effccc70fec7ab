// Micro-benchmarks of a few instruction throughputs on sm_100a (issue cycles per warp-instruction per SM sub-partition),
// used to choose between sort designs (scripts/ubench/README in DESIGN.md section 4.1).  Not part of the library.
#include <cstdio>
#include <cuda_runtime.h>
#define N_IT 4096
template <int OP>
__global__ void k(unsigned* out, long long* cyc, unsigned seed) {
    __shared__ unsigned hist[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    unsigned x = seed * (threadIdx.x + 1) * 2654435761u, acc = 0;
    float f = (float)(threadIdx.x + 1) * 1.37f;
    double d = 0.0;
    long long t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < N_IT; i++) {
        x = x * 1664525u + 1013904223u;
        if (OP == 0) acc += x >> 24;                                                   // baseline: LCG + shift + add
        if (OP == 1) acc += __match_any_sync(0xffffffffu, x >> 24);                     // 8-bit digit, ~uniform
        if (OP == 2) acc += __match_any_sync(0xffffffffu, (x >> 30));                   // 4 distinct values
        if (OP == 3) acc += atomicAdd(&hist[x >> 24], 1u);                             // smem atomic, uniform bins
        if (OP == 4) acc += atomicAdd(&hist[x >> 30], 1u);                             // smem atomic, 4 bins (conflicts)
        if (OP == 5) { f = f + __uint_as_float((x >> 9) | 0x3f800000u); d += (double)f; } // F2F.F64.F32 + DADD
        if (OP == 6) { d = d + __hiloint2double(0x3ff00000 | (x >> 12), x); }         // DADD only
        if (OP == 7) acc += __shfl_xor_sync(0xffffffffu, x, 1);                        // SHFL
        if (OP == 8) { f = fminf(f, __uint_as_float((x >> 9) | 0x3f800000u)); }        // FMNMX
        if (OP == 9) acc += __popc(x) + __ffs(x);
        if (OP == 10) { hist[(threadIdx.x * 33 + (x >> 27)) & 1023] = x; acc += hist[(threadIdx.x * 33 + i) & 1023]; }  // STS + LDS
        if (OP == 11) { const double dd = (double)__uint_as_float((x >> 9) | 0x3f800000u); d = __fma_rn(dd, dd, d); }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + (unsigned)f + (unsigned)d + hist[threadIdx.x];
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP>
void run(const char* name, int warps_per_sm) {
    unsigned* out; long long* cyc;
    const int blocks = 148, threads = warps_per_sm * 32;
    cudaMalloc(&out, blocks * threads * 4); cudaMalloc(&cyc, blocks * 8);
    k<OP><<<blocks, threads>>>(out, cyc, 12345u);
    k<OP><<<blocks, threads>>>(out, cyc, 12345u);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double m = 0; for (int i = 0; i < blocks; i++) m += h[i]; m /= blocks;
    // cycles per loop iteration per warp slot on one sub-partition: total cycles / (iterations * warps per SMSP)
    printf("%-28s warps/SM %2d  cycles/iter/warp-per-SMSP %.2f\n", name, warps_per_sm, m / N_IT / (warps_per_sm / 4.0));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int w : {16, 32}) {
        if (w == 16) {
            run<0>("baseline lcg", 16); run<1>("match_any 8-bit uniform", 16); run<2>("match_any 2-bit", 16);
            run<3>("smem atomicAdd 256 bins", 16); run<4>("smem atomicAdd 4 bins", 16); run<5>("F2F.F64.F32 + FADD + DADD", 16);
            run<6>("DADD", 16); run<7>("SHFL", 16); run<8>("FMNMX", 16); run<9>("POPC+FFS", 16); run<10>("STS+LDS", 16);
            run<11>("F2F + DFMA", 16);
        } else {
            run<0>("baseline lcg", 32); run<1>("match_any 8-bit uniform", 32); run<3>("smem atomicAdd 256 bins", 32);
            run<5>("F2F.F64.F32 + FADD + DADD", 32); run<6>("DADD", 32);
        }
    }
    return 0;
}
