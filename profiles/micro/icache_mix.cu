// Microbenchmark: instruction SUPPLY rate for code that does not fit the L0 instruction cache (B200, sm_100a).
// Straight-line body of N DFMAs, each followed by F independent 1-cycle integer instructions: the FP64 pipe
// needs one DFMA per 2 cycles per SMSP, so with F fillers the front end must deliver (1 + F) / 2 instructions
// per cycle per SMSP to keep it busy. Small bodies (fit L0) are looped for comparison.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o icache_mix icache_mix.cu && ./icache_mix
#include <cstdio>
#include <cuda_runtime.h>

template <int N, int F>
__global__ void __launch_bounds__(128, 4) k(double* out, unsigned* iout, int reps, double m, double c, unsigned z)
{
    double a[8];
    unsigned u[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x + i;
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = threadIdx.x * 7 + i;
#pragma unroll 1
    for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int i = 0; i < N / 8; ++i) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                a[kk] = fma(a[kk], m, c + (double)(i & 7));
#pragma unroll
                for (int f = 0; f < F; ++f) u[(kk + f) & 3] = (u[(kk + f) & 3] ^ z) + (unsigned)(i * 8 + kk + 13 * f);
            }
        }
    }
    double s = 0; unsigned v = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) v ^= u[i];
    if (s == 123.456) out[0] = s;
    if (v == 0x12345u) iout[0] = v;
}

template <int N, int F> void run(int reps)
{
    double* out; unsigned* iout; cudaMalloc(&out, 8); cudaMalloc(&iout, 4);
    const int grid = 148 * 4 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<N, F><<<grid, 128>>>(out, iout, reps, 1.0000001, 1e-9, 0x9e3779b9u);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<N, F><<<grid, 128>>>(out, iout, reps, 1.0000001, 1e-9, 0x9e3779b9u);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double dfma = (double)grid * 4 * N * reps;
    const double cyc = ms * 1e-3 * 1.965e9 * 148 * 4;
    printf("body %5d DFMA + %d fillers each, reps %3d: FP64 pipe %.1f%%, ~%.2f inst/clk/SMSP if every filler is one instruction\n", N, F, reps,
           100.0 * dfma * 2.0 / cyc, dfma * (1 + F) / cyc);
    cudaFree(out); cudaFree(iout);
}

int main()
{
    run<128, 0>(64); run<128, 1>(64); run<128, 2>(64); run<128, 3>(64);        // small bodies (L0 / L1.5 resident), looped
    run<4096, 0>(2); run<4096, 1>(2); run<4096, 2>(2); run<4096, 3>(2);        // 64 KB+ of straight-line code
    run<1024, 1>(8); run<1024, 2>(8); run<2048, 1>(4); run<2048, 2>(4);
    return 0;
}
