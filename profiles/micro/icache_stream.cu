// Microbenchmark: cost of STRAIGHT-LINE code that exceeds the instruction caches on B200 (sm_100a).
// A kernel body of N independent-chain DFMAs (no loop inside the body) is executed `reps` times per thread by
// 16 warps per SM; reported: FP64 pipe utilisation vs code size. Tells how large a scene-specialised
// (fully unrolled) kernel may grow before instruction fetch, not the FP64 pipe, sets the pace.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o icache_stream icache_stream.cu && ./icache_stream
#include <cstdio>
#include <cuda_runtime.h>

template <int N> struct Body {
    __device__ __forceinline__ static void run(double (&a)[8], double m, double c)
    {
#pragma unroll
        for (int i = 0; i < N / 8; ++i) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fma(a[k], m, c + (double)(i & 7));   // distinct immediates keep ptxas from re-rolling
        }
    }
};

template <int N>
__global__ void __launch_bounds__(128, 4) k(double* out, int reps, double m, double c)
{
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x + i;
#pragma unroll 1
    for (int r = 0; r < reps; ++r) Body<N>::run(a, m, c);
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 123.456) out[0] = s;
}

template <int N> void run(int reps, int blocks_per_sm)
{
    double* out; cudaMalloc(&out, 8);
    const int grid = 148 * blocks_per_sm * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<N><<<grid, 128>>>(out, reps, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<N><<<grid, 128>>>(out, reps, 1.0000001, 1e-9);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double inst = (double)grid * 4 * N * reps;                    // warp DFMAs
    const double cyc = ms * 1e-3 * 1.965e9 * 148 * 4;                   // SMSP cycles at max clock
    printf("body %6d DFMA (%4d KB) reps %d: %.3f ms, FP64 pipe %.1f%% of 2 cyc/warp-inst\n", N, N * 16 / 1024, reps, ms, 100.0 * inst * 2.0 / cyc);
    cudaFree(out);
}

int main()
{
    run<256>(64, 4); run<1024>(16, 4); run<2048>(8, 4); run<4096>(4, 4); run<8192>(2, 4); run<16384>(1, 4);
    run<2048>(1, 4); run<4096>(1, 4); run<8192>(1, 4);
    return 0;
}
