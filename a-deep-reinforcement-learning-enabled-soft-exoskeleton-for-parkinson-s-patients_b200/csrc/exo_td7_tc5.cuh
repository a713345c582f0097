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

#ifdef TC5_TIMELINE
__device__ long long tc5_marks[64];
#define TC5_MARK(i) do { if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (i) < 64) tc5_marks[i] = clock64(); } while (0)
#else
#define TC5_MARK(i) do { } while (0)
#endif

constexpr int TM5 = 128, KA = 32;                   // tile rows, k per atom (one 128-byte swizzle row)
constexpr int PACK_WARPS = 8, THREADS5 = (PACK_WARPS + 1) * 32;   // 8 packing / epilogue warps + the MMA-issuing warp

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}

template <int TN, int NS>
struct Smem5 {
    static constexpr int A_BYTES = TM5 * 128, B_BYTES = TN * 128;             // one atom of A / B (hi or lo)
    static constexpr int STAGE = 2 * A_BYTES + 2 * B_BYTES;                   // A_hi | A_lo | B_hi | B_lo
    static constexpr int EPI_PITCH = TN + 1;
    static constexpr int BYTES = NS * STAGE + 1024 /* alignment slack */ + 128 /* barriers, tmem slot */;
    static_assert(NS * STAGE >= TM5 * EPI_PITCH * 4, "the epilogue tile reuses the operand stages");
};

template <bool TA, bool TB, int TN, int NS>
__global__ void __launch_bounds__(THREADS5, NS == 2 ? 2 : 1) gemm_tc5_kernel(GemmArgs g) {
    using S = Smem5<TN, NS>;
    extern __shared__ unsigned char smem_raw5[];
    unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw5) + 1023) & ~(uintptr_t)1023);   // 1024-aligned atoms
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + NS * S::STAGE);       // full[NS] | free[NS] | accumulator done
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NS + 1);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int m0 = blockIdx.y * TM5, n0 = blockIdx.x * TN;
    const int kbeg = blockIdx.z * g.kchunk, kend = min(g.K, kbeg + g.kchunk);
    const int n_atoms = (kend - kbeg + KA - 1) / KA;
    if (n_atoms <= 0) return;
    const uint32_t bar0 = smem_u32(bars);
    auto bar_full = [&](int s) { return bar0 + 8u * s; };
    auto bar_free = [&](int s) { return bar0 + 8u * (NS + s); };
    const uint32_t bar_done = bar0 + 8u * (2 * NS);
    TC5_MARK(0);

    // Thread -> element map of one atom for the 8 packing warps: 16 A elements and TN / 8 B elements per thread, rows
    // row0 + 8 u (u = 0 .. 15 / TN/8 - 1) at one k of the atom, so that a thread's shared-memory offsets are op_off(row0, kk)
    // + 1024 u (one 8-row group per u) and its global addresses one pointer + u * stride. Chosen so that a warp's loads cover
    // whole 32-byte sectors AND its stores fall on 32 different banks of the swizzled image:
    //   operand contiguous along k: the warp takes row (warp + 8 u), lanes = the 32 k      (bank = ((k / 4) ^ row) * 4 + k % 4)
    //   operand contiguous along m / n: the warp takes k-group `warp` (4 k) of rows 8 u .. 8 u + 7, lane = (row % 8, k % 4)
    constexpr int NA = TM5 / 8, NB = TN / 8;
    const int ml = lane & 7, kl = lane >> 3;
    const int a_row0 = TA ? ml : warp, a_kk = TA ? warp * 4 + kl : lane;
    const int b_row0 = TB ? warp : ml, b_kk = TB ? lane : warp * 4 + kl;
    const uint32_t a_off = op_off(a_row0, a_kk), b_off = op_off(b_row0, b_kk);
    const float* pa = TA ? g.A + (size_t)(kbeg + a_kk) * g.lda + (m0 + a_row0) : g.A + (size_t)(m0 + a_row0) * g.lda + (kbeg + a_kk);
    const float* pb = TB ? g.B + (size_t)(n0 + b_row0) * g.ldb + (kbeg + b_kk) : g.B + (size_t)(kbeg + b_kk) * g.ldb + (n0 + b_row0);
    const int a_stride = TA ? 8 : 8 * g.lda, b_stride = TB ? 8 * g.ldb : 8;
    const size_t a_step = TA ? (size_t)KA * g.lda : (size_t)KA, b_step = TB ? (size_t)KA : (size_t)KA * g.ldb;
    const int a_ulim = max(g.M - m0 - a_row0 + 7, 0) >> 3, b_ulim = max(g.N - n0 - b_row0 + 7, 0) >> 3;   // rows row0 + 8 u inside the matrix
    float ra[NA], rb[NB];
    auto load_atom = [&](int atom) {
        const int k0 = kbeg + atom * KA;
        const int au = k0 + a_kk < kend ? a_ulim : 0, bu = k0 + b_kk < kend ? b_ulim : 0;
#pragma unroll
        for (int u = 0; u < NA; ++u) ra[u] = u < au ? __ldg(pa + u * a_stride) : 0.f;
#pragma unroll
        for (int u = 0; u < NB; ++u) rb[u] = u < bu ? __ldg(pb + u * b_stride) : 0.f;
        pa += a_step; pb += b_step;
    };

    if (warp < PACK_WARPS) load_atom(0);              // in flight across the set-up below
    if (t == 0) {
#pragma unroll
        for (int s = 0; s < NS; ++s) { mbar_init(bar_full(s), PACK_WARPS * 32); mbar_init(bar_free(s), 1); }
        mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == PACK_WARPS) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    constexpr uint32_t IDESC = idesc_tf32(TN, TM5);
    TC5_MARK(1);

    if (warp == PACK_WARPS) {
        // ---- the MMA warp: per atom 4 k-steps x 3 products into the one accumulator, small terms first
        int stage = 0, phase = 0;
        for (int atom = 0; atom < n_atoms; ++atom) {
            mbar_wait(bar_full(stage), phase);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t ah = smem_u32(base + stage * S::STAGE), al = ah + S::A_BYTES, bh = al + S::A_BYTES, bl = bh + S::B_BYTES;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {     // 8 k per MMA = 32 bytes along the swizzled row
                    const uint32_t o = ks * 32;
                    tc_mma_tf32(tmem, smem_desc(al + o), smem_desc(bh + o), IDESC, (atom | ks) != 0);
                    tc_mma_tf32(tmem, smem_desc(ah + o), smem_desc(bl + o), IDESC, 1);
                    tc_mma_tf32(tmem, smem_desc(ah + o), smem_desc(bh + o), IDESC, 1);
                }
                tc_commit(bar_free(stage));          // arrives when these MMAs have read the stage
                if (atom == n_atoms - 1) tc_commit(bar_done);
            }
            __syncwarp();
            if (++stage == NS) { stage = 0; phase ^= 1; }
        }
    } else {
        // ---- the packing warps
        int stage = 0, phase = 1;                    // parity of the stage's PREVIOUS use; first NS atoms do not wait
        for (int atom = 0; atom < n_atoms; ++atom) {
            if (atom >= NS) mbar_wait(bar_free(stage), phase);
            TC5_MARK(8 + atom * 4);
            float* dah = reinterpret_cast<float*>(base + stage * S::STAGE + a_off);
            float* dbh = reinterpret_cast<float*>(base + stage * S::STAGE + 2 * S::A_BYTES + b_off);
#pragma unroll
            for (int u = 0; u < NA; ++u) {
                const float x = ra[u], h = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
                dah[u * 256] = h; dah[u * 256 + S::A_BYTES / 4] = x - h;
            }
#pragma unroll
            for (int u = 0; u < NB; ++u) {
                const float x = rb[u], h = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
                dbh[u * 256] = h; dbh[u * 256 + S::B_BYTES / 4] = x - h;
            }
            fence_proxy_async_smem();                // generic-proxy writes -> visible to the tensor core's async proxy
            mbar_arrive(bar_full(stage));
            TC5_MARK(9 + atom * 4);
            if (atom + 1 < n_atoms) load_atom(atom + 1);
            TC5_MARK(10 + atom * 4);
            if (++stage == NS) { stage = 0; phase ^= 1; }
        }
    }

    // ---- epilogue: TMEM -> registers -> shared (row-major tile), then rows with the lanes along n (coalesced), loads batched
    mbar_wait(bar_done, 0);
    tc_fence_after();
    TC5_MARK(3);
    float* tile = reinterpret_cast<float*>(base);    // [128][TN + 1]; every MMA has finished reading the stages
    if (warp < PACK_WARPS) {
        const int q = warp & 3, half = warp >> 2, r = q * 32 + lane;       // TMEM lane quarter of this warp, column half
        constexpr int HC = TN / 2;                    // 32 or 40 columns
#pragma unroll
        for (int c0 = 0; c0 < HC; c0 += 8) {
            uint32_t v[8];
            tmem_ld8(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * HC + c0), v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 8; ++j) tile[r * S::EPI_PITCH + half * HC + c0 + j] = __uint_as_float(v[j]);
        }
    }
    tc_fence_before();
    __syncthreads();
    TC5_MARK(4);
    if (warp == PACK_WARPS) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
        return;
    }
    constexpr int NJ = (TN + 31) / 32, RB = 4;        // columns per lane, rows per batch
    float bj[NJ];
    bool cv[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
        const int c = lane + 32 * j;
        cv[j] = c < TN && n0 + c < g.N;
        bj[j] = (g.bias && cv[j]) ? __ldg(g.bias + n0 + c) : 0.f;
    }
    const int act = g.act, dact = g.dact, out = g.out;
    for (int r0 = warp * (TM5 / PACK_WARPS); r0 < (warp + 1) * (TM5 / PACK_WARPS); r0 += RB) {
        float d[RB][NJ], old[RB][NJ];
#pragma unroll
        for (int i = 0; i < RB; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int m = m0 + r0 + i, n = n0 + lane + 32 * j;
                const bool ok = cv[j] && m < g.M;
                d[i][j] = (g.D && ok) ? __ldg(g.D + (size_t)m * g.ldd + n) : 0.f;
                old[i][j] = (out == OUT_ADD && ok) ? g.C[(size_t)m * g.ldc + n] : 0.f;
            }
#pragma unroll
        for (int i = 0; i < RB; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int m = m0 + r0 + i, c = lane + 32 * j;
                if (!(cv[j] && m < g.M)) continue;
                float v = act_fwd(tile[(r0 + i) * S::EPI_PITCH + c] + bj[j], act);
                if (g.D) v *= act_bwd(d[i][j], dact);
                float* cp = g.C + (size_t)m * g.ldc + n0 + c;
                if (out == OUT_STORE) *cp = v;
                else if (out == OUT_ADD) *cp = old[i][j] + v;
                else atomicAdd(cp, v);
            }
    }
    TC5_MARK(5);
}

template <bool TA, bool TB>
inline cudaError_t launch(const GemmArgs& g, int tn, int ns, dim3 grid, cudaStream_t s) {
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(gemm_tc5_kernel<TA, TB, 64, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem5<64, 2>::BYTES);
        cudaFuncSetAttribute(gemm_tc5_kernel<TA, TB, 80, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem5<80, 2>::BYTES);
        cudaFuncSetAttribute(gemm_tc5_kernel<TA, TB, 64, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem5<64, 3>::BYTES);
        cudaFuncSetAttribute(gemm_tc5_kernel<TA, TB, 80, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem5<80, 3>::BYTES);
        configured = true;
    }
    if (ns == 2) {
        if (tn == 64) gemm_tc5_kernel<TA, TB, 64, 2><<<grid, THREADS5, Smem5<64, 2>::BYTES, s>>>(g);
        else gemm_tc5_kernel<TA, TB, 80, 2><<<grid, THREADS5, Smem5<80, 2>::BYTES, s>>>(g);
    } else {
        if (tn == 64) gemm_tc5_kernel<TA, TB, 64, 3><<<grid, THREADS5, Smem5<64, 3>::BYTES, s>>>(g);
        else gemm_tc5_kernel<TA, TB, 80, 3><<<grid, THREADS5, Smem5<80, 3>::BYTES, s>>>(g);
    }
    return cudaGetLastError();
}

}  // namespace tc5
