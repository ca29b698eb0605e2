// ubench_ilp.cu -- FFMA issue rate per SMSP as a function of resident warps per SMSP and of the
// number of independent FMA chains per thread (ILP).  sm_100a.  Answers: how many warps does
// the pair kernel need to keep the FMA pipes busy when its per-thread ILP is low?
#include <cuda_runtime.h>
#include <stdio.h>

template <int ILP>
__global__ void __launch_bounds__(1024) k(float* sink, int iters) {
    float f[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) f[i] = threadIdx.x * 1e-3f + i;
    float m = 0.999f + blockIdx.x * 1e-9f, c = 1e-3f + threadIdx.x * 1e-9f;
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 64 / ILP; ++u) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(m), "f"(c));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += f[i];
    if (s == 123.456f) *sink = s;
}

template <int ILP>
void run(float* sink, int sms, double hz) {
    for (int wps : {1, 2, 3, 4, 6, 8}) {
        const int threads = 128 * wps;  // one block per SM, wps warps per SMSP
        const int iters = 20000;
        k<ILP><<<sms, threads>>>(sink, 100);
        cudaDeviceSynchronize();
        cudaEvent_t a, b;
        cudaEventCreate(&a); cudaEventCreate(&b);
        float best = 1e30f;
        for (int r = 0; r < 3; ++r) {
            cudaEventRecord(a);
            k<ILP><<<sms, threads>>>(sink, iters);
            cudaEventRecord(b);
            cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (ms < best) best = ms;
        }
        const double winstr = (double)sms * (threads / 32) * iters * 64.0;
        printf("{\"ilp\": %d, \"warps_per_smsp\": %d, \"ffma_per_cycle_per_smsp\": %.3f}\n", ILP, wps,
               winstr / (best * 1e-3) / (sms * 4.0 * hz));
    }
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    float* sink;
    cudaMalloc(&sink, 4);
    for (int r = 0; r < 300; ++r) k<8><<<prop.multiProcessorCount, 1024>>>(sink, 2000);
    cudaDeviceSynchronize();
    run<1>(sink, prop.multiProcessorCount, khz * 1e3);
    run<2>(sink, prop.multiProcessorCount, khz * 1e3);
    run<4>(sink, prop.multiProcessorCount, khz * 1e3);
    run<8>(sink, prop.multiProcessorCount, khz * 1e3);
    return 0;
}
