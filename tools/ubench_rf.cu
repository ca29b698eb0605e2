// ubench_rf.cu -- FFMA issue rate versus the number of register-file reads per instruction
// (sm_100a).  d[i] = fma(x, y, d[i]) with x / y either one loop-invariant register (served by the
// operand-reuse cache) or a different register per instruction.
#include <cuda_runtime.h>
#include <stdio.h>

template <int FRESH>
__global__ void __launch_bounds__(512) k(float* sink, int iters) {
    float a[8], b[8], d[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = 0.999f + i * 1e-6f + threadIdx.x * 1e-9f; b[i] = 1e-3f + i * 1e-6f; d[i] = i + blockIdx.x; }
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (FRESH == 1) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(d[i]) : "f"(a[0]), "f"(b[0]));
                if (FRESH == 2) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(d[i]) : "f"(a[i]), "f"(b[0]));
                if (FRESH == 3) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(d[i]) : "f"(a[i]), "f"(b[(i + u) & 7]));
                if (FRESH == 4) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(d[i]) : "f"(a[(i + u) & 7]), "f"(b[(i + 3 * u + 1) & 7]));
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += d[i] + a[i] + b[i];
    if (s == 123.456f) *sink = s;
}

template <int FRESH>
void run(float* sink, int sms, double hz) {
    for (int wps : {2, 3, 4}) {
        const int threads = 128 * wps, iters = 20000;
        k<FRESH><<<sms, threads>>>(sink, 100);
        cudaDeviceSynchronize();
        cudaEvent_t a, b;
        cudaEventCreate(&a); cudaEventCreate(&b);
        float best = 1e30f;
        for (int r = 0; r < 3; ++r) {
            cudaEventRecord(a);
            k<FRESH><<<sms, threads>>>(sink, iters);
            cudaEventRecord(b);
            cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (ms < best) best = ms;
        }
        printf("{\"fresh_operands\": %d, \"warps_per_smsp\": %d, \"ffma_per_cycle_per_smsp\": %.3f}\n", FRESH, wps,
               (double)sms * (threads / 32) * iters * 64.0 / (best * 1e-3) / (sms * 4.0 * hz));
    }
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    float* sink;
    cudaMalloc(&sink, 4);
    for (int r = 0; r < 300; ++r) k<1><<<prop.multiProcessorCount, 512>>>(sink, 4000);
    cudaDeviceSynchronize();
    run<1>(sink, prop.multiProcessorCount, khz * 1e3);
    run<2>(sink, prop.multiProcessorCount, khz * 1e3);
    run<3>(sink, prop.multiProcessorCount, khz * 1e3);
    run<4>(sink, prop.multiProcessorCount, khz * 1e3);
    return 0;
}
