// ubench_tmem.cu -- can tensor memory serve as a per-thread scratchpad (sm_100a)?
// Each thread owns one TMEM lane (32x32b shape: lane i of warp w <-> TMEM lane 32*(w%4)+i) and
// keeps NCOL fp32 "accumulators" in consecutive columns.  The kernel does read-modify-write passes
// over them with tcgen05.ld / tcgen05.st (x8 = 8 columns per instruction), checks the result and
// reports the sustained RMW bandwidth per SM.  Used to decide where the pair kernel of D = 13..16
// keeps its K x (1 + D + D(D+1)/2) statistics.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "f"(v[0]),
                 "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]));
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int NCOL, int ALLOC>
__global__ void __launch_bounds__(128) k(int iters, int* errors, long long* cycles) {
    __shared__ uint32_t base_s;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(&base_s)),
                     "n"(ALLOC));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t base = base_s + ((uint32_t)(warp & 3) << 21);  // lane field = bits 31:16, 32 lanes per warp
    float v[8];
#pragma unroll
    for (int c = 0; c < NCOL; c += 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
        tmem_st8(base + c, v);
    }
    tmem_wait_st();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < NCOL; c += 8) {
            tmem_ld8(base + c, v);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] += (float)(threadIdx.x + c + i + 1);
            tmem_st8(base + c, v);
        }
        tmem_wait_st();
    }
    const long long t1 = clock64();
    int bad = 0;
#pragma unroll
    for (int c = 0; c < NCOL; c += 8) {
        tmem_ld8(base + c, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 8; ++i) bad += (v[i] != (float)iters * (float)(threadIdx.x + c + i + 1));
    }
    if (bad) atomicAdd(errors, bad);
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base_s), "n"(ALLOC));
}

template <int NCOL, int ALLOC>
void run(int sms, int ctas_per_sm) {
    int* err;
    long long* cyc;
    cudaMalloc(&err, 4);
    cudaMemset(err, 0, 4);
    cudaMalloc(&cyc, 8 * sms * ctas_per_sm);
    const int iters = 2000;
    k<NCOL, ALLOC><<<sms * ctas_per_sm, 128>>>(10, err, cyc);
    cudaMemset(err, 0, 4);
    k<NCOL, ALLOC><<<sms * ctas_per_sm, 128>>>(iters, err, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    int herr = -1;
    long long hc = 0;
    cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    // bytes read (= bytes written) per SM per cycle
    const double bytes = (double)ctas_per_sm * 128 * NCOL * 4.0 * iters;
    printf("{\"ncol\": %d, \"alloc\": %d, \"ctas_per_sm\": %d, \"cuda\": \"%s\", \"errors\": %d, \"cycles\": %lld, "
           "\"rmw_read_bytes_per_cycle_per_sm\": %.1f}\n",
           NCOL, ALLOC, ctas_per_sm, cudaGetErrorString(e), herr, hc, bytes / (double)hc);
    cudaFree(err);
    cudaFree(cyc);
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    run<64, 64>(prop.multiProcessorCount, 1);
    run<64, 64>(prop.multiProcessorCount, 4);
    run<152, 256>(prop.multiProcessorCount, 1);
    run<152, 256>(prop.multiProcessorCount, 2);
    return 0;
}
