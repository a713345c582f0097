// mma_env.cu - what slows a stream of tcgen05.mma k-steps (M=128, tf32, N=256 then N=48 on one A operand) down:
// flags bit0: 4 warps spin on an mbarrier (try_wait) meanwhile; bit1: A operand with 9728-byte 8-row-group stride;
// bit2: a producer thread streams cp.async.bulk copies (9728 B per k-step) into shared memory meanwhile;
// bit3: tcgen05.commit after every k-step; bit4: tcgen05.fence::after_thread_sync before every k-step.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done;
}
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t sbo) {
    return (uint64_t)((addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | (6ull << 61);
}
__device__ __forceinline__ void commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
constexpr int XB = 155648, ST = 9728;
__global__ void k(int flags, int reps, const unsigned char* src, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bars[8];
    __shared__ uint32_t tslot;
    __shared__ volatile int stop;
    for (int i = threadIdx.x; i < 220 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 8; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        stop = 0;
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
    const int warp = threadIdx.x >> 5;
    const uint32_t done_bar = smem_u32(&bars[0]), never_bar = smem_u32(&bars[1]), dummy_bar = smem_u32(&bars[2]), tma_bar = smem_u32(&bars[3]);
    if (threadIdx.x == 0) {
        const uint32_t ID256 = (1u << 4) | (2u << 7) | (2u << 10) | (32u << 17) | (8u << 24), ID48 = (ID256 & ~(63u << 17)) | (6u << 17);
        const uint32_t a0 = smem_u32(smem), b0 = a0 + XB;
        const uint32_t asbo = (flags & 2) ? 9728u : 256u;
        const long long t0 = clock64();
#pragma unroll 1
        for (int i = 0; i < reps; ++i) {
            if (flags & 16) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t ks = (uint32_t)(i % 38);
            const uint64_t ad = desc(a0 + ks * 256u, asbo);
            const uint32_t w = b0 + (uint32_t)(i % 6) * ST;
            mma(tmem, ad, desc(w, 256), ID256, i > 0);
            mma(tmem + 256, ad, desc(w + 8192, 256), ID48, i > 0);
            if (flags & 8) commit(dummy_bar);
        }
        const long long t1 = clock64();
        commit(done_bar);
        while (!mbar_try(done_bar, 0)) {}
        const long long t2 = clock64();
        out[0] = t1 - t0; out[1] = t2 - t0;
        stop = 1;
    } else if (warp == 1 && (threadIdx.x & 31) == 0 && (flags & 4)) {
        // bulk-copy stream into the 6 weight stages, as fast as the engine takes them (no consumer handshake)
        uint32_t par = 0;
        long long n = 0;
        while (!stop) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tma_bar), "r"(6u * ST) : "memory");
            for (int s = 0; s < 6; ++s)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem) + XB + s * ST),
                             "l"(src + (size_t)((n * 6 + s) % 200) * ST), "r"((uint32_t)ST), "r"(tma_bar) : "memory");
            while (!mbar_try(tma_bar, par)) {}
            par ^= 1; ++n;
        }
        out[2] = n * 6;
    } else if (warp >= 2 && (flags & 1)) {
        while (!stop) { mbar_try(never_bar, 0); }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}
int main() {
    long long* d; cudaMalloc(&d, 64);
    unsigned char* src; cudaMalloc(&src, 200 * ST); cudaMemset(src, 0, 200 * ST);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int reps = 380;
    printf("flags(spin=1,bigSBO=2,tma=4,commit=8,fence=16) issue_cyc_per_kstep total_cyc_per_kstep tma_stages\n");
    const int fl[] = {0, 1, 2, 4, 8, 16, 3, 6, 7, 15, 31};
    for (int f : fl) {
        cudaMemset(d, 0, 64);
        k<<<1, 192, 220 * 1024>>>(f, reps, src, d);
        long long h[3];
        cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        printf("%d %.1f %.1f %lld\n", f, (double)h[0] / reps, (double)h[1] / reps, h[2]);
    }
    return 0;
}
