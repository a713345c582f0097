// exo_td7_tc5.cuh — the TD7 update's GEMM on the 5th-generation tensor cores: tcgen05.mma kind::tf32 with the accumulator
// in tensor memory, 3xTF32 (fp32-class products, as the mma.sync kernel it replaces).
//
//   C[M x N] (+)= op(A)[M x K] op(B)[K x N],  same GemmArgs / epilogue (bias, activation, derivative, store / add / atomic)
//   as gemm_kernel in exo_td7.cu: forward Y = act(X W^T + b), input gradient dX = (dY W) * act'(out), weight gradient
//   dW += dY^T X (split over the batch).
//
// One CTA = one 128 x TN output tile (UMMA M = 128 = the 128 TMEM lanes; TN = 64 or 80 accumulator columns), 128 threads:
//   * all four warps PACK the operands: a k-atom of 32 columns of the A tile (128 rows) and of the B tile (TN rows) is read
//     from global memory with the lanes along the operand's contiguous axis (whichever of the four transposition cases),
//     every value is split in registers into hi = tf32(x) and lo = x - hi, and both halves are written into the canonical
//     K-major SWIZZLE_128B shared-memory image the tensor core reads (rows 128 B apart inside an 8-row group, groups 1024 B
//     apart, 16-byte chunk c of row r at position c ^ (r & 7)). Row pitches of 87 / 307 / 920 floats rule out TMA boxes and
//     16-byte copies, which is why this is register packing and not cp.async.bulk.tensor;
//   * one elected thread issues, per 8-wide k-step, three MMAs into the same accumulator: lo_A hi_B, hi_A lo_B, hi_A hi_B
//     (small terms first), then tcgen05.commit on the stage's mbarrier; two stages, so atom k+1 is packed while atom k
//     multiplies;
//   * epilogue: each warp reads its 32 accumulator rows from TMEM (tcgen05.ld 32x32b), the tile crosses to row-major through
//     shared memory, and the threads apply gemm_store with the lanes along n (coalesced).
// The 1920-row batch gives 15 row tiles; with TN = 80 a 300-wide layer is 60 CTAs, a weight gradient (300 x 300 outputs,
// K = 1920 split 8 ways) 96 CTAs.
#pragma once

namespace tc5 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
// A wait that never returns would hang the GPU box: after ~2 s of polling, trap instead.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (uint32_t spins = 0; !done; ++spins) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && (spins & 1023u) == 1023u) {
            const long long t = clock64();
            if (t0 == 0) t0 = t;
            else if (t - t0 > 4000000000ll) __trap();
        }
    }
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// SWIZZLE_128B K-major operand: start >> 4, LBO (unused inside one swizzle atom) = 1, SBO = 1024 B between 8-row groups,
// descriptor version 1 (sm_100), layout type 2. (Same encoding as csrc/exo_policy.cu, where it is measured.)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
    return (uint64_t)((addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// byte offset of element (row r, k in [0, 32)) of a one-atom operand tile
__device__ __forceinline__ uint32_t op_off(uint32_t r, uint32_t k) {
    return (r >> 3) * 1024u + (r & 7u) * 128u + ((((k >> 2) & 7u) ^ (r & 7u)) << 4) + (k & 3u) * 4u;
}
// D fp32, A and B tf32, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr uint32_t idesc_tf32(int n, int m) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void split_store(unsigned char* hi, unsigned char* lo, uint32_t off, float x) {
    const float h = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
    *reinterpret_cast<float*>(hi + off) = h;
    *reinterpret_cast<float*>(lo + off) = x - h;
}

constexpr int TM5 = 128, KA = 32;                   // tile rows, k per atom (one 128-byte swizzle row)

template <int TN>
struct Smem5 {
    static constexpr int A_BYTES = TM5 * 128, B_BYTES = TN * 128;             // one atom of A / B (hi or lo)
    static constexpr int STAGE = 2 * A_BYTES + 2 * B_BYTES;                   // A_hi | A_lo | B_hi | B_lo
    static constexpr int EPI_PITCH = TN + 1;
    static constexpr int BYTES = 2 * STAGE + 1024 /* alignment slack */ + 64 /* barriers, tmem slot */;
    static_assert(2 * STAGE >= TM5 * EPI_PITCH * 4, "the epilogue tile reuses the operand stages");
};

template <bool TA, bool TB, int TN>
__global__ void __launch_bounds__(128) gemm_tc5_kernel(GemmArgs g) {
    using S = Smem5<TN>;
    extern __shared__ unsigned char smem_raw5[];
    unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw5) + 1023) & ~(uintptr_t)1023);   // 1024-aligned atoms
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + 2 * S::STAGE);        // [0,1] stage free, [2] accumulator done
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int m0 = blockIdx.y * TM5, n0 = blockIdx.x * TN;
    const int kbeg = blockIdx.z * g.kchunk, kend = min(g.K, kbeg + g.kchunk);
    if (t == 0) {
        mbar_init(smem_u32(&bars[0]), 1); mbar_init(smem_u32(&bars[1]), 1); mbar_init(smem_u32(&bars[2]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    constexpr uint32_t IDESC = idesc_tf32(TN, TM5);

    const int n_atoms = (kend - kbeg + KA - 1) / KA;
    for (int atom = 0; atom < n_atoms; ++atom) {
        const int stage = atom & 1, k0 = kbeg + atom * KA;
        unsigned char* sAh = base + stage * S::STAGE;
        unsigned char* sAl = sAh + S::A_BYTES;
        unsigned char* sBh = sAl + S::A_BYTES;
        unsigned char* sBl = sBh + S::B_BYTES;
        if (atom >= 2) mbar_wait(smem_u32(&bars[stage]), ((atom >> 1) - 1) & 1);      // the MMAs that read this stage two atoms ago
        // ---- pack A: 128 rows x 32 k
        if (!TA) {                                   // A[m * lda + k]: lanes along k, a warp walks 32 rows
#pragma unroll 4
            for (int rr = 0; rr < 32; ++rr) {
                const int r = warp * 32 + rr, gm = m0 + r, gk = k0 + lane;
                const float x = (gm < g.M && gk < kend) ? g.A[(size_t)gm * g.lda + gk] : 0.f;
                split_store(sAh, sAl, op_off(r, lane), x);
            }
        } else {                                     // A[k * lda + m]: lanes along m, a warp walks the 32 k of its 32 rows
            const int r = warp * 32 + lane, gm = m0 + r;
#pragma unroll 4
            for (int kk = 0; kk < 32; ++kk) {
                const int gk = k0 + kk;
                const float x = (gm < g.M && gk < kend) ? g.A[(size_t)gk * g.lda + gm] : 0.f;
                split_store(sAh, sAl, op_off(r, kk), x);
            }
        }
        // ---- pack B: TN rows (n) x 32 k
        if (TB) {                                    // B[n * ldb + k]: lanes along k
            for (int r = warp; r < TN; r += 4) {
                const int gn = n0 + r, gk = k0 + lane;
                const float x = (gn < g.N && gk < kend) ? g.B[(size_t)gn * g.ldb + gk] : 0.f;
                split_store(sBh, sBl, op_off(r, lane), x);
            }
        } else {                                     // B[k * ldb + n]: lanes along n
            for (int e = t; e < TN * 32; e += 128) {
                const int r = e % TN, kk = e / TN;
                const int gn = n0 + r, gk = k0 + kk;
                const float x = (gn < g.N && gk < kend) ? g.B[(size_t)gk * g.ldb + gn] : 0.f;
                split_store(sBh, sBl, op_off(r, kk), x);
            }
        }
        fence_proxy_async_smem();                    // generic-proxy writes -> visible to the tensor core's async proxy
        __syncthreads();
        if (t == 0) {
            tc_fence_after();
            const uint32_t ah = smem_u32(sAh), al = smem_u32(sAl), bh = smem_u32(sBh), bl = smem_u32(sBl);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {         // 8 k per MMA = 32 bytes along the swizzled row
                const uint32_t o = ks * 32;
                tc_mma_tf32(tmem, smem_desc(al + o), smem_desc(bh + o), IDESC, (atom | ks) != 0);
                tc_mma_tf32(tmem, smem_desc(ah + o), smem_desc(bl + o), IDESC, 1);
                tc_mma_tf32(tmem, smem_desc(ah + o), smem_desc(bh + o), IDESC, 1);
            }
            tc_commit(smem_u32(&bars[stage]));       // arrives when these MMAs have read the stage
            if (atom == n_atoms - 1) tc_commit(smem_u32(&bars[2]));
        }
    }
    // ---- epilogue: TMEM -> registers -> shared (row-major tile) -> gemm_store with the lanes along n
    mbar_wait(smem_u32(&bars[2]), 0);
    tc_fence_after();
    float* tile = reinterpret_cast<float*>(base);    // [128][TN + 1]; every MMA has finished reading the stages
    {
        const int r = warp * 32 + lane;
#pragma unroll
        for (int c0 = 0; c0 < TN; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) tile[r * S::EPI_PITCH + c0 + j] = __uint_as_float(v[j]);
        }
    }
    tc_fence_before();
    __syncthreads();
    for (int e = t; e < TM5 * TN; e += 128) {
        const int r = e / TN, c = e - r * TN;
        gemm_store(g, m0 + r, n0 + c, tile[r * S::EPI_PITCH + c]);
    }
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
    }
}

template <bool TA, bool TB>
inline cudaError_t launch(const GemmArgs& g, int tn, dim3 grid, cudaStream_t s) {
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(gemm_tc5_kernel<TA, TB, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem5<64>::BYTES);
        cudaFuncSetAttribute(gemm_tc5_kernel<TA, TB, 80>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem5<80>::BYTES);
        configured = true;
    }
    if (tn == 64) gemm_tc5_kernel<TA, TB, 64><<<grid, 128, Smem5<64>::BYTES, s>>>(g);
    else gemm_tc5_kernel<TA, TB, 80><<<grid, 128, Smem5<80>::BYTES, s>>>(g);
    return cudaGetLastError();
}

}  // namespace tc5
