// mma_fetch.cu - cycles per k-step (tf32, M=128, N=160 then N=144 on one A slice, SS operands) when every MMA reads
// FRESH shared-memory operands (A walks 38 k-slices of a 152 KB activation tile, B walks 6-7 weight stages), for the
// K-major SWIZZLE_32B layout of csrc/exo_policy.cu (mode 0) and for SWIZZLE_128B (mode 1: 128 B rows, 4 k-steps per
// swizzle atom). mode 2: SWIZZLE_32B with only the N=160 MMA; mode 3: same operands every time (no fresh fetch).
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
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, %1;\n\tselp.b32 %0, 1, 0, px;\n\t}" : "=r"(pred) : "r"(0xFFFFFFFFu));
    return pred != 0;
}
__host__ __device__ constexpr uint32_t idesc(int n) { return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24); }
__global__ void k(int mode, int reps, int fill, int spin, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar, never;
    __shared__ volatile int stop;
    __shared__ uint32_t tslot;
    for (int i = threadIdx.x; i < 224 * 1024 / 4; i += blockDim.x) {
        uint32_t h = (uint32_t)i * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        ((float*)smem)[i] = fill ? ((float)(h & 0xFFFF) / 65536.0f - 0.5f) : 0.0f;      // zeros vs values in (-0.5, 0.5): is the rate data dependent?
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&never)));
        stop = 0;
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
    if (threadIdx.x < 32) {
        const uint32_t base = smem_u32(smem);
        const bool sw128 = mode == 1;
        const uint32_t a_sbo = sw128 ? 1024u : 9728u, b_sbo = sw128 ? 1024u : 256u, layout = sw128 ? 2u : 6u;
        const uint32_t wbase = base + 155648 + (sw128 ? 8192 : 0);      // SW128: keep 1024 B alignment
        const long long t0 = clock64();
#pragma unroll 1
        for (int rep = 0; rep < reps; ++rep) {
#pragma unroll 1
            for (int ks = 0; ks < 38; ks += 2) {
                if (elect_one()) {
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int kk = mode == 3 ? 0 : ks + u;
                        uint32_t a_addr, b_addr, b_hi;
                        if (sw128) {   // A: [k-atom (4 k-steps)][16 row groups][8 rows x 128 B]; B: one k-atom of 304 rows = 38912 B (fits once)
                            a_addr = base + (uint32_t)(kk >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
                            b_addr = wbase + (uint32_t)(kk & 3) * 32u;
                            b_hi = b_addr + 160u * 128u;
                        } else {
                            a_addr = base + (uint32_t)kk * 256u;
                            b_addr = wbase + (uint32_t)(kk % 6) * 9728u;
                            b_hi = b_addr + 20u * 256u;
                        }
                        const uint64_t ad = desc(a_addr, a_sbo, layout);
                        const uint32_t acc = (rep | ks | u) != 0;
                        mma(tmem, ad, desc(b_addr, b_sbo, layout), idesc(160), acc);
                        if (mode != 2) mma(tmem + 160, ad, desc(b_hi, b_sbo, layout), idesc(144), acc);
                    }
                }
                __syncwarp();
            }
        }
        const long long t1 = clock64();
        if (elect_one()) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        __syncwarp();
        while (!mbar_try(smem_u32(&bar), 0)) {}
        const long long t2 = clock64();
        if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; stop = 1; }
    } else if (spin) {
        // the epilogue warps of the real kernel: polling an mbarrier while the MMAs run
        while (!stop) mbar_try(smem_u32(&never), 0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}
int main() {
    long long* d; cudaMalloc(&d, 64);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
    const int reps = 10;
    printf("mode(0 sw32, 1 sw128, 2 sw32 N=160 only, 3 sw32 same operands) issue_cyc_per_kstep total_cyc_per_kstep\n");
    for (int spin = 0; spin < 2; ++spin)
    for (int mode = 0; mode < 4; ++mode) {
        const int fill = 1;
        k<<<1, 320, 224 * 1024>>>(mode, reps, fill, spin, d);
        long long h[2];
        cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        printf("spin=%d %d %.1f %.1f\n", spin, mode, (double)h[0] / (reps * 38), (double)h[1] / (reps * 38));
    }
    return 0;
}
