// mma_rate.cu - cycles per tcgen05.mma (cta_group::1, M=128, SS operands) as a function of N, kind and smem layout.
// One CTA; one thread issues `reps` MMAs back to back on fixed shared-memory operands, commits, waits. Tuning aid for
// csrc/exo_policy.cu (numbers quoted in DESIGN.md).   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
template <int KIND>   // 0 tf32, 1 f16(bf16)
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    if (KIND == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) |
           (1ull << 46) | ((uint64_t)layout << 61);
}

// layout: 0 none (LBO 128, SBO 256), 6 SW32 (SBO 256), 2 SW128 (SBO 1024)
template <int KIND>
__global__ void k(int n, int layout, int reps, int nacc, int adv, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tslot;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tslot;
    if (threadIdx.x == 0) {
        const uint32_t fmt = KIND == 0 ? 2u : 1u;
        const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24);
        const uint32_t sbo = layout == 2 ? 1024 : 256, lbo = layout == 0 ? 128 : 16;
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
        uint32_t par = 0;
        for (int rep = 0; rep < 3; ++rep) {
            const long long t0 = clock64();
            const uint64_t ad = desc(a0, lbo, sbo, layout), bd = desc(b0, lbo, sbo, layout);
            const uint64_t step = (uint64_t)(adv >> 4);
            const uint32_t d1 = tmem + (nacc > 1 ? (uint32_t)n : 0u);
#pragma unroll 1
            for (int i = 0; i < reps; i += 8) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    mma<KIND>((j & 1) ? d1 : tmem, ad + j * step, bd + j * step, idesc, (i + j) >= 2);
            }
            const long long t1 = clock64();
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
            mbar_wait(smem_u32(&bar), par); par ^= 1;
            const long long t2 = clock64();
            out[rep * 2] = t1 - t0; out[rep * 2 + 1] = t2 - t0;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

int main() {
    long long* d; cudaMalloc(&d, 64);
    cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int reps = 256;
    const int ns[] = {16, 48, 64, 128, 160, 256};
    const int layouts[] = {0, 6, 2};
    printf("kind layout N nacc adv issue_cyc_per_mma total_cyc_per_mma\n");
    for (int kind = 0; kind < 2; ++kind)
        for (int layout : layouts)
            for (int n : ns)
                for (int nacc = 1; nacc <= 2; ++nacc)
                    for (int adv = 0; adv <= 256; adv += 256) {
                        if (nacc * n > 512) continue;
                        if (layout == 2 && adv) continue;
                        if (kind == 0) k<0><<<1, 128, 200 * 1024>>>(n, layout, reps, nacc, adv, d);
                        else k<1><<<1, 128, 200 * 1024>>>(n, layout, reps, nacc, adv, d);
                        long long h[6];
                        cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
                        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                        printf("%s %d %d %d %d %.1f %.1f\n", kind ? "bf16" : "tf32", layout, n, nacc, adv, (double)h[4] / reps, (double)h[5] / reps);
                    }
    return 0;
}
