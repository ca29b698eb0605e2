// ubench_f32x2.cu -- issue-slot microbenchmarks for the pair-kernel design (sm_100a).
//
// Question: the pair kernel is issue-bound (about half of its instructions are not FMAs).
// Does the packed FFMA2 (fma.rn.f32x2) free issue slots, i.e. can LDS / MUFU / ALU / SHFL
// instructions be issued "under" a stream of FFMA2 without slowing it down?
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_f32x2 ubench_f32x2.cu
// Prints one JSON line per test: warp-instructions / cycle / SMSP (assuming the SM clock
// reported by the device) and FP32 TFLOP/s.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CHECK(x)                                                                        \
    do {                                                                                \
        cudaError_t e = (x);                                                            \
        if (e != cudaSuccess) {                                                         \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            return 1;                                                                   \
        }                                                                               \
    } while (0)

__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t fmul2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float ffma(float a, float b, float c) {
    float d;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float fadd(float a, float b) {
    float d;
    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ float ex2(float a) {
    float d;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a));
    return d;
}
__device__ __forceinline__ float lds(uint32_t addr) {
    float d;
    asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(d) : "r"(addr));
    return d;
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 d;
    asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w) : "r"(addr));
    return d;
}
__device__ __forceinline__ uint32_t iadd(uint32_t a, uint32_t b) {
    uint32_t d;
    asm volatile("add.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ float shfl(float a) {
    float d;
    asm volatile("shfl.sync.bfly.b32 %0, %1, 1, 0x1f, 0xffffffff;" : "=f"(d) : "f"(a));
    return d;
}

enum {
    T_FFMA = 0,       // 8 scalar FFMA chains
    T_FFMA2,          // 8 FFMA2 chains
    T_FMUL2,
    T_FADD2,
    T_FFMA_LDS,       // 8 FFMA + 8 LDS.32
    T_FFMA2_LDS,      // 8 FFMA2 + 8 LDS.32
    T_FFMA2_LDS128,   // 8 FFMA2 + 8 LDS.128 (broadcast address)
    T_FFMA_MUFU,      // 8 FFMA + 2 MUFU
    T_FFMA2_MUFU,     // 8 FFMA2 + 2 MUFU
    T_FFMA2_MUFU4,    // 8 FFMA2 + 4 MUFU
    T_FFMA_FADD,      // 8 FFMA + 8 FADD (both fma-type)
    T_FFMA2_FADD,     // 8 FFMA2 + 8 scalar FADD
    T_FFMA_IADD,      // 8 FFMA + 8 IADD (alu pipe)
    T_FFMA2_IADD,     // 8 FFMA2 + 8 IADD
    T_FFMA2_IADD16,   // 8 FFMA2 + 16 IADD
    T_FFMA_SHFL,      // 8 FFMA + 2 SHFL
    T_FFMA2_SHFL,     // 8 FFMA2 + 2 SHFL
    T_FFMA2_FFMA,     // 8 FFMA2 + 8 scalar FFMA
    T_FFMA2_MIX,      // 8 FFMA2 + 4 LDS + 2 MUFU + 4 IADD + 4 FADD
    T_FFMA_MIX,       // 16 FFMA + 4 LDS + 2 MUFU + 4 IADD + 4 FADD   (same math, scalar)
    T_FFMA16_LDS4,    // 16 FFMA + 4 LDS.32
    T_FFMA2_LDS4,     // 8 FFMA2 + 4 LDS.32
    T_FFMA16_MUFU1,   // 16 FFMA + 1 MUFU
    T_FFMA2_MUFU1,    // 8 FFMA2 + 1 MUFU
    T_FFMA2_4CH,      // 4 FFMA2 chains only (ILP 4)
    T_FFMA2_DISTINCT, // 8 FFMA2, distinct b/c operands per chain
    T_COUNT
};

const char* kNames[T_COUNT] = {"ffma", "ffma2", "fmul2", "fadd2", "ffma+lds", "ffma2+lds", "ffma2+lds128",
                               "ffma+mufu(8:2)", "ffma2+mufu(8:2)", "ffma2+mufu(8:4)", "ffma+fadd", "ffma2+fadd",
                               "ffma+iadd", "ffma2+iadd", "ffma2+iadd(8:16)", "ffma+shfl(8:2)", "ffma2+shfl(8:2)",
                               "ffma2+ffma", "ffma2+mix", "ffma(16)+mix", "ffma(16)+lds(4)", "ffma2(8)+lds(4)", "ffma(16)+mufu(1)", "ffma2(8)+mufu(1)", "ffma2 x4 chains", "ffma2 distinct operands"};
// warp-instructions per loop body, fp32 flops per thread per loop body
const int kInstr[T_COUNT] = {8, 8, 8, 8, 16, 16, 16, 10, 10, 12, 16, 16, 16, 16, 24, 10, 10, 16, 22, 30};
const int kFlops[T_COUNT] = {16, 32, 16, 16, 16, 32, 32, 16, 32, 32, 24, 40, 16, 32, 32, 16, 32, 48, 36, 36, 32, 32, 32, 32, 16, 32};

template <int T>
__global__ void __launch_bounds__(256) bench_kernel(float* sink, int iters) {
    __shared__ float sh[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sh[i] = i * 1e-6f;
    __syncthreads();
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sh);
    const uint32_t my = sbase + 4u * threadIdx.x;           // conflict-free
    const uint32_t bc = sbase + 16u * (threadIdx.x >> 5);   // broadcast within the warp

    float f[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = threadIdx.x * 1e-3f + i;
    uint64_t p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float2 v = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f + i);
        p[i] = *reinterpret_cast<uint64_t*>(&v);
    }
    float2 mv = make_float2(0.999f + blockIdx.x * 1e-9f, 0.998f);
    float2 cv = make_float2(1e-3f, 2e-3f);
    const uint64_t m2 = *reinterpret_cast<uint64_t*>(&mv);
    const uint64_t c2 = *reinterpret_cast<uint64_t*>(&cv);
    const float m = mv.x, c = cv.x;
    float e0 = 0.1f, e1 = 0.2f, e2 = 0.3f, e3 = 0.4f;
    float g[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = 0.01f * i;
    uint32_t n[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) n[i] = threadIdx.x + i;
    float s0 = 1.f, s1 = 2.f;
    float ldacc = 0.f;

#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (T == T_FFMA) {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = ffma(f[i], m, c);
            } else if (T == T_FFMA2) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], m2, c2);
            } else if (T == T_FMUL2) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = fmul2(p[i], m2);
            } else if (T == T_FADD2) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = fadd2(p[i], c2);
            } else if (T == T_FFMA_LDS) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { f[i] = ffma(f[i], m, c); ldacc = lds(my + 1024u * (i & 3)); }
            } else if (T == T_FFMA2_LDS) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { p[i] = ffma2(p[i], m2, c2); ldacc = lds(my + 1024u * (i & 3)); }
            } else if (T == T_FFMA2_LDS128) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { p[i] = ffma2(p[i], m2, c2); float4 v = lds128(bc + 256u * (i & 3)); ldacc = v.x; }
            } else if (T == T_FFMA_MUFU) {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = ffma(f[i], m, c);
                e0 = ex2(e0); e1 = ex2(e1);
            } else if (T == T_FFMA2_MUFU) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], m2, c2);
                e0 = ex2(e0); e1 = ex2(e1);
            } else if (T == T_FFMA2_MUFU4) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], m2, c2);
                e0 = ex2(e0); e1 = ex2(e1); e2 = ex2(e2); e3 = ex2(e3);
            } else if (T == T_FFMA_FADD) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { f[i] = ffma(f[i], m, c); g[i] = fadd(g[i], c); }
            } else if (T == T_FFMA2_FADD) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { p[i] = ffma2(p[i], m2, c2); g[i] = fadd(g[i], c); }
            } else if (T == T_FFMA_IADD) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { f[i] = ffma(f[i], m, c); n[i] = iadd(n[i], n[(i + 1) & 7]); }
            } else if (T == T_FFMA2_IADD) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { p[i] = ffma2(p[i], m2, c2); n[i] = iadd(n[i], n[(i + 1) & 7]); }
            } else if (T == T_FFMA2_IADD16) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    p[i] = ffma2(p[i], m2, c2);
                    n[i] = iadd(n[i], n[(i + 1) & 7]);
                    n[8 + i] = iadd(n[8 + i], n[8 + ((i + 1) & 7)]);
                }
            } else if (T == T_FFMA_SHFL) {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = ffma(f[i], m, c);
                s0 = shfl(s0); s1 = shfl(s1);
            } else if (T == T_FFMA2_SHFL) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], m2, c2);
                s0 = shfl(s0); s1 = shfl(s1);
            } else if (T == T_FFMA2_FFMA) {
#pragma unroll
                for (int i = 0; i < 8; ++i) { p[i] = ffma2(p[i], m2, c2); f[i] = ffma(f[i], m, c); }
            } else if (T == T_FFMA2_MIX) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    p[i] = ffma2(p[i], m2, c2);
                    if (i & 1) ldacc = lds(my + 1024u * (i >> 1));
                    else n[i] = iadd(n[i], n[i + 1]);
                    if (i < 4) g[i] = fadd(g[i], c);
                }
                e0 = ex2(e0); e1 = ex2(e1);
            } else if (T == T_FFMA16_LDS4) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    f[i] = ffma(f[i], m, c);
                    f[8 + i] = ffma(f[8 + i], m, c);
                    if (i & 1) ldacc = lds(my + 1024u * (i >> 1));
                }
            } else if (T == T_FFMA2_LDS4) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    p[i] = ffma2(p[i], m2, c2);
                    if (i & 1) ldacc = lds(my + 1024u * (i >> 1));
                }
            } else if (T == T_FFMA16_MUFU1) {
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = ffma(f[i], m, c);
                e0 = ex2(e0);
            } else if (T == T_FFMA2_MUFU1) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], m2, c2);
                e0 = ex2(e0);
            } else if (T == T_FFMA2_4CH) {
#pragma unroll
                for (int i = 0; i < 4; ++i) p[i] = ffma2(p[i], m2, c2);
            } else if (T == T_FFMA2_DISTINCT) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], p[(i + 3) & 7], p[(i + 5) & 7]);
            } else if (T == T_FFMA_MIX) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    f[i] = ffma(f[i], m, c);
                    f[8 + i] = ffma(f[8 + i], m, c);
                    if (i & 1) ldacc = lds(my + 1024u * (i >> 1));
                    else n[i] = iadd(n[i], n[i + 1]);
                    if (i < 4) g[i] = fadd(g[i], c);
                }
                e0 = ex2(e0); e1 = ex2(e1);
            }
        }
    }
    float s = ldacc + e0 + e1 + e2 + e3 + s0 + s1;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += f[i] + (float)n[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float2 v = *reinterpret_cast<float2*>(&p[i]);
        s += v.x + v.y + g[i];
    }
    if (s == 123.456f) *sink = s;
}

template <int T>
int run(float* sink, int sms, double clock_hz, int blocks_per_sm, int threads) {
    const int iters = 16384;
    const int blocks = sms * blocks_per_sm;
    bench_kernel<T><<<blocks, threads>>>(sink, 64);
    CHECK(cudaDeviceSynchronize());
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(a);
        bench_kernel<T><<<blocks, threads>>>(sink, iters);
        cudaEventRecord(b);
        CHECK(cudaEventSynchronize(b));
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    const double warps = (double)blocks * threads / 32.0;
    const double winstr = warps * iters * 4.0 * kInstr[T];
    const double flops = (double)blocks * threads * iters * 4.0 * kFlops[T];
    const double ipc = winstr / (best * 1e-3) / (sms * 4.0 * clock_hz);
    printf("{\"test\": \"%s\", \"warps_per_smsp\": %d, \"ms\": %.4f, \"warp_instr_per_cycle_per_smsp\": %.3f, \"tflops\": %.2f}\n",
           kNames[T], blocks_per_sm * threads / 128, best, ipc, flops / (best * 1e-3) / 1e12);
    fflush(stdout);
    return 0;
}

template <int T>
int run_all(float* sink, int sms, double hz) {
    if (run<T>(sink, sms, hz, 2, 256)) return 1;   // 4 warps / SMSP
    if (run<T>(sink, sms, hz, 4, 256)) return 1;   // 8 warps / SMSP
    if constexpr (T + 1 < T_COUNT) return run_all<T + 1>(sink, sms, hz);
    return 0;
}

int main() {
    cudaDeviceProp prop;
    CHECK(cudaGetDeviceProperties(&prop, 0));
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", prop.name, prop.multiProcessorCount, khz);
    float* sink;
    CHECK(cudaMalloc(&sink, 4));
    for (int r = 0; r < 400; ++r) bench_kernel<T_FFMA><<<prop.multiProcessorCount * 4, 256>>>(sink, 8192);
    CHECK(cudaDeviceSynchronize());
    return run_all<0>(sink, prop.multiProcessorCount, khz * 1e3);
}
