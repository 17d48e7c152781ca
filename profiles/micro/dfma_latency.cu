// Microbenchmark: FP64 pipe issue/latency behaviour on B200 -- how much ILP x warps saturates the DFMA pipe.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dfma_latency dfma_latency.cu && ./dfma_latency
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void k(double* out, int iters, double m, double c, long long* cyc)
{
    double a[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) a[i] = threadIdx.x + i;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u)
#pragma unroll
            for (int i = 0; i < ILP; ++i) a[i] = fma(a[i], m, c);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += a[i];
    if (s == 123.456) out[0] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int ILP> void run(int warps_per_smsp)
{
    double* out; long long* cyc; cudaMalloc(&out, 8); cudaMalloc(&cyc, 8);
    const int iters = 2000;
    const int threads = 32 * 4 * warps_per_smsp;   // one block per SM, 4 SMSPs
    k<ILP><<<148, threads>>>(out, iters, 1.0000001, 1e-9, cyc);
    cudaDeviceSynchronize();
    k<ILP><<<148, threads>>>(out, iters, 1.0000001, 1e-9, cyc);
    cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double n_inst = (double)iters * 16 * ILP;             // per warp
    const double per_smsp = n_inst * warps_per_smsp;            // warp-instructions per SMSP
    printf("ILP %d warps/SMSP %d: %.2f cycles per dependent step, pipe busy %.1f%% (2 cycles per warp DFMA)\n", ILP, warps_per_smsp,
           (double)h / (iters * 16.0), 100.0 * per_smsp * 2.0 / (double)h);
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    for (int w : {1, 2, 3, 4, 5, 6, 8}) { run<1>(w); run<2>(w); run<3>(w); run<4>(w); run<8>(w); }
    return 0;
}
