// exo_td7_tc5.cuh — the TD7 update's GEMM on the 5th-generation tensor cores: tcgen05.mma kind::tf32 with the accumulator
// in tensor memory, 3xTF32 (fp32-class products, as the mma.sync kernel it replaces as the default).
//
//   C[M x N] (+)= op(A)[M x K] op(B)[K x N],  same GemmArgs / epilogue semantics (bias, activation, derivative, store / add /
//   atomic) as gemm_kernel in exo_td7.cu: forward Y = act(X W^T + b), input gradient dX = (dY W) * act'(out), weight
//   gradient dW += dY^T X (split over the batch).
//
// One CTA = one 128 x TN output tile (UMMA M = 128 = the 128 TMEM lanes; TN = 64 or 80 accumulator columns), 9 warps:
//   * warps 0-7 PACK the operands: a k-atom of 32 columns of the A tile (128 rows) and of the B tile (TN rows) is read from
//     global memory (predicated 4-byte loads: row pitches of 87 / 307 floats rule out TMA boxes and 16-byte copies), every
//     value is split in registers into hi = tf32(x) and lo = x - hi, and both halves are written into the canonical K-major
//     SWIZZLE_128B shared-memory image the tensor core reads (rows 128 B apart inside an 8-row group, groups 1024 B apart,
//     16-byte chunk c of row r at position c ^ (r & 7)). The thread -> element maps make every warp load whole 32-byte
//     sectors and every warp store 32 different banks, for all four transposition cases. Two atoms live in registers, so
//     the loads of atom k + 1 are in flight while atom k is packed; the stage is handed over with fence.proxy.async +
//     mbarrier.arrive on full[stage];
//   * warp 8 allocates the accumulator and issues, per 8-wide k-step, three MMAs into it: lo_A hi_B, hi_A lo_B, hi_A hi_B
//     (small terms first), then tcgen05.commit on free[stage]; two stages (52 KB each at TN = 80: the image is written at
//     the shared-memory store rate, 416 cycles per atom, next to ~600-900 cycles of MMA issue per atom);
//   * epilogue: TMEM -> registers -> row-major tile in shared memory, then rows with the lanes along n (coalesced).
// Optional split-k over a thread-block cluster (CK = 2 / 4) with the partial tiles summed through distributed shared memory;
// see the kernel. Measured alone (tools/ubench/tc5_gemm.cu, 1920 x 300 x 300 forward with ELU, B200): 5.4 -> 1.0 ms per TD7
// iteration came from the pipeline (first version: one warp group doing load / pack / issue in sequence), 25 -> 14 us per
// launch from the epilogue (see there); the whole TD7 iteration is 0.73 ms against 1.18 ms for mma.sync.
#pragma once
#include <type_traits>

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

// x = ok ? *p : 0 as one predicated load (written in PTX so that the address arithmetic stays outside the predicate: the
// compiler's own form of `ok ? __ldg(p) : 0.f` re-derived every address under its predicate, 8.6 instructions per load)
__device__ __forceinline__ float ldg_if(const float* p, bool ok) {
    float x;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\tmov.f32 %0, 0f00000000;\n\t@p ld.global.nc.f32 %0, [%1];\n\t}"
                 : "=f"(x) : "l"(p), "r"((int)ok));
    return x;
}

// Shared memory through its own window (32-bit addresses): the tile pointer below comes out of an integer round-up, after
// which the compiler no longer knows the address space and emits generic LD / ST.
__device__ __forceinline__ void sts_f32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ float lds_f32(uint32_t a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }

// ELU's negative branch, exp(x) - 1, in ~13 instructions: a degree-6 Taylor polynomial above -0.25 (next term 5e-8 relative)
// and ex2.approx(x log2 e) - 1 below (|result| > 0.22, relative error < 5e-7); expm1f is ~4x that and, at two warps per
// scheduler, was a third of this kernel.
__device__ __forceinline__ float elu_fast(float x) {
    float p = fmaf(x, 1.f / 720.f, 1.f / 120.f);
    p = fmaf(p, x, 1.f / 24.f); p = fmaf(p, x, 1.f / 6.f); p = fmaf(p, x, 0.5f); p = fmaf(p, x, 1.f);
    const float small = p * x, big = __expf(x) - 1.f;
    return x > 0.f ? x : (x > -0.25f ? small : big);
}

template <int TN, int NS>
struct Smem5 {
    static constexpr int A_BYTES = TM5 * 128, B_BYTES = TN * 128;             // one atom of A / B (hi or lo)
    static constexpr int STAGE = 2 * A_BYTES + 2 * B_BYTES;                   // A_hi | A_lo | B_hi | B_lo
    static constexpr int EPI_PITCH = TN + 1;
    static constexpr int BYTES = NS * STAGE + 1024 /* alignment slack */ + 128 /* barriers, tmem slot */;
    static_assert((2 * NS + 2) * 8 + 4 <= 128, "barriers + tmem slot");
    static_assert(NS * STAGE >= TM5 * EPI_PITCH * 4, "the epilogue tile reuses the operand stages");
};

// thread-block cluster helpers (split-k across CK CTAs, partial tiles summed through distributed shared memory)
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
__device__ __forceinline__ float lds_cluster_f32(uint32_t a) { float v; asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }

// CK = 1: one CTA per (tile, k-chunk blockIdx.z), as gemm_kernel. CK = 2 / 4 (store / add outputs): a cluster of CK CTAs along
// z shares one output tile, CTA `rank` multiplies the rank-th part of k, every CTA leaves its partial tile in its own shared
// memory, and after one cluster barrier CTA `rank` finishes rows [rank 128/CK, (rank + 1) 128/CK): own partial + the peers'
// through ld.shared::cluster, then bias / activation / derivative / store. A 1920-row batch is only 15 row tiles: without
// this a 300-wide layer occupies 60 of the 148 SMs for the whole k loop.
template <bool TA, bool TB, int TN, int NS, int CK>
__global__ void __launch_bounds__(THREADS5, NS == 2 ? 2 : 1) gemm_tc5_kernel(GemmArgs g) {
    using S = Smem5<TN, NS>;
    extern __shared__ unsigned char smem_raw5[];
    unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw5) + 1023) & ~(uintptr_t)1023);   // 1024-aligned atoms
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + NS * S::STAGE);       // full[NS] | free[NS] | accumulator done | tmem ready
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NS + 2);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int m0 = blockIdx.y * TM5, n0 = blockIdx.x * TN;
    const int crank = CK > 1 ? (int)cluster_ctarank() : 0;
    int kbeg, kend;
    if (CK == 1) { kbeg = blockIdx.z * g.kchunk; kend = min(g.K, kbeg + g.kchunk); }
    else {                                            // (the host launches CK > 1 only when every rank gets at least one atom)
        const int per = ((g.K + KA - 1) / KA + CK - 1) / CK * KA;
        kbeg = crank * per; kend = min(g.K, kbeg + per);
    }
    const int n_atoms = (kend - kbeg + KA - 1) / KA;
    if (CK == 1 && n_atoms <= 0) return;
    const uint32_t bar0 = smem_u32(bars);
    auto bar_full = [&](int s) { return bar0 + 8u * s; };
    auto bar_free = [&](int s) { return bar0 + 8u * (NS + s); };
    const uint32_t bar_done = bar0 + 8u * (2 * NS), bar_tmem = bar0 + 8u * (2 * NS + 1);
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
    float ra0[NA], rb0[NB], ra1[NA], rb1[NB];        // two atoms in registers: atom k + 1 is in flight while atom k is packed
    auto load_atom = [&](int atom, float* ra, float* rb) {
        const int k0 = kbeg + atom * KA;
        const int au = k0 + a_kk < kend ? a_ulim : 0, bu = k0 + b_kk < kend ? b_ulim : 0;
        const float* p = pa;
#pragma unroll
        for (int u = 0; u < NA; ++u) { ra[u] = ldg_if(p, u < au); p += a_stride; }
        p = pb;
#pragma unroll
        for (int u = 0; u < NB; ++u) { rb[u] = ldg_if(p, u < bu); p += b_stride; }
        pa += a_step; pb += b_step;
    };

    if (warp < PACK_WARPS) load_atom(0, ra0, rb0);   // in flight across the set-up below
    // the epilogue's per-lane constants: column validity and bias of columns lane + 32 j (loaded now, used after the k loop)
    constexpr int NJ = (TN + 31) / 32;
    bool cv[NJ];
    float bj[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
        cv[j] = lane + 32 * j < TN && n0 + lane + 32 * j < g.N;
        bj[j] = ldg_if(g.bias + n0 + lane + 32 * j, g.bias != nullptr && cv[j]);
    }
    if (t == 0) {
#pragma unroll
        for (int s = 0; s < NS; ++s) { mbar_init(bar_full(s), PACK_WARPS * 32); mbar_init(bar_free(s), 1); }
        mbar_init(bar_done, 1); mbar_init(bar_tmem, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    constexpr uint32_t IDESC = idesc_tf32(TN, TM5);
    TC5_MARK(1);

    if (warp == PACK_WARPS) {
        // ---- the MMA warp: allocates the accumulator (128 TMEM columns; the packing warps do not wait for this), then per
        // atom 4 k-steps x 3 products into the one accumulator, small terms first
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_tmem);
        mbar_wait(bar_tmem, 0);
        tc_fence_after();
        const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
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
        const uint32_t base_s = smem_u32(base);
        auto pack_atom = [&](int atom, const float* ra, const float* rb) {
            if (atom >= NS) mbar_wait(bar_free(stage), phase);
            TC5_MARK(8 + atom * 4);
            const uint32_t dah = base_s + stage * S::STAGE + a_off, dbh = base_s + stage * S::STAGE + 2 * S::A_BYTES + b_off;
#pragma unroll
            for (int u = 0; u < NA; ++u) {
                const float x = ra[u], h = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
                sts_f32(dah + u * 1024, h); sts_f32(dah + u * 1024 + S::A_BYTES, x - h);
            }
#pragma unroll
            for (int u = 0; u < NB; ++u) {
                const float x = rb[u], h = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
                sts_f32(dbh + u * 1024, h); sts_f32(dbh + u * 1024 + S::B_BYTES, x - h);
            }
            fence_proxy_async_smem();                // generic-proxy writes -> visible to the tensor core's async proxy
            mbar_arrive(bar_full(stage));
            TC5_MARK(9 + atom * 4);
            if (++stage == NS) { stage = 0; phase ^= 1; }
        };
        for (int atom = 0; atom < n_atoms; atom += 2) {
            if (atom + 1 < n_atoms) load_atom(atom + 1, ra1, rb1);
            pack_atom(atom, ra0, rb0);
            if (atom + 1 >= n_atoms) break;
            if (atom + 2 < n_atoms) load_atom(atom + 2, ra0, rb0);
            pack_atom(atom + 1, ra1, rb1);
        }
    }

    // ---- epilogue. This kernel runs 2 warps per scheduler from a cold instruction cache, so the epilogue is written for few
    // instructions and independent chains (the first version - 4 rows x 3 columns unrolled around gemm_store - was 645
    // instructions per batch and 40 % of the kernel; a rolled one-value-at-a-time activation loop was 250 cycles per element
    // of dependent-issue latency):
    //   1. each packing warp reads its TMEM lane quarter / column half 8 columns at a time and writes the row-major tile
    //      of raw sums to shared memory;
    //   2. [CK > 1] cluster barrier; from here a CTA owns 128 / CK rows and adds the peers' partial tiles;
    //   3. each warp finishes its rows with the lanes along n (coalesced), 4 rows per batch, every load of a batch issued
    //      before the first use: [+ peers] + bias, activation, multiply by act'(saved output) - one branch-free form for all
    //      activations - then store / add / atomic. Complete batches take a path without row predicates.
    const uint32_t tile = smem_u32(base);            // [128][TN + 1] floats, over the operand stages
    const int act = g.act, dact = g.dact, out = g.out;
    if (n_atoms > 0) { mbar_wait(bar_done, 0); mbar_wait(bar_tmem, 0); }     // every MMA has finished reading the stages
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
    TC5_MARK(3);
    if (warp < PACK_WARPS) {
        const int q = warp & 3, half = warp >> 2, r = q * 32 + lane;       // TMEM lane quarter of this warp, column half
        constexpr int HC = TN / 2;                    // 32 or 40 columns
        const int cend = min((half + 1) * HC, g.N - n0);                    // columns beyond N are never read back
        const uint32_t trow = tile + (uint32_t)(r * S::EPI_PITCH) * 4u;
#pragma unroll 1
        for (int c0 = half * HC; c0 < cend; c0 += 8) {
            uint32_t vi[8];
            if (n_atoms > 0) tmem_ld8(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, vi);
            else {
#pragma unroll
                for (int j = 0; j < 8; ++j) vi[j] = 0u;
            }
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 8; ++j) sts_f32(trow + (uint32_t)(c0 + j) * 4u, __uint_as_float(vi[j]));
        }
    }
    tc_fence_before();
    __syncthreads();
    TC5_MARK(4);
    if (CK > 1) cluster_sync_all();                  // every partial tile of the cluster is in shared memory
    if (warp == PACK_WARPS) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
        if (CK == 1) return;
    }
    if (warp < PACK_WARPS) {
        constexpr int ROWS = TM5 / PACK_WARPS / CK, RB = ROWS < 4 ? ROWS : 4;
        const int rbeg = crank * (TM5 / CK) + warp * ROWS, rend = min(rbeg + ROWS, g.M - m0);
        const bool has_d = g.D != nullptr, is_add = out == OUT_ADD, is_atomic = out == OUT_ATOMIC;
        // act'(y) from the saved OUTPUT y, branch-free: y > thr ? 1 : a0 + a1 y + a2 y^2
        //   none: 1 | relu: y > 0 ? 1 : 0 | elu: y > 0 ? 1 : y + 1 | tanh: 1 - y^2
        const float thr = (dact == ACT_RELU || dact == ACT_ELU) ? 0.f : __int_as_float(0x7f800000);
        const float a0 = dact == ACT_RELU ? 0.f : 1.f, a1 = dact == ACT_ELU ? 1.f : 0.f, a2 = dact == ACT_TANH ? -1.f : 0.f;
        const float* dp = g.D + (size_t)(m0 + rbeg) * g.ldd + (n0 + lane);
        float* cp = g.C + (size_t)(m0 + rbeg) * g.ldc + (n0 + lane);
        uint32_t tp = tile + (uint32_t)(rbeg * S::EPI_PITCH + lane) * 4u;
        uint32_t peer[CK > 1 ? CK - 1 : 1];
#pragma unroll
        for (int pi = 0; pi < CK - 1; ++pi) peer[pi] = mapa_u32(tp, (uint32_t)((crank + 1 + pi) % CK));
        TC5_MARK(6);
        // one batch of RB rows; FULL (a std::true_type / false_type value): all RB rows are inside the matrix
        auto batch = [&](auto full, int nrow) {
            constexpr bool FULL = decltype(full)::value;
            float d[RB][NJ], old[RB][NJ], v[RB][NJ];
            if (has_d) {
                const float* dpi = dp;
#pragma unroll
                for (int i = 0; i < RB; ++i) {
#pragma unroll
                    for (int j = 0; j < NJ; ++j) d[i][j] = ldg_if(dpi + 32 * j, cv[j] && (FULL || i < nrow));
                    dpi += g.ldd;
                }
            }
            if (is_add) {
                const float* cpi = cp;
#pragma unroll
                for (int i = 0; i < RB; ++i) {
#pragma unroll
                    for (int j = 0; j < NJ; ++j) old[i][j] = ldg_if(cpi + 32 * j, cv[j] && (FULL || i < nrow));   // (read and written by this thread only)
                    cpi += g.ldc;
                }
            }
#pragma unroll
            for (int i = 0; i < RB; ++i)             // (the volatile shared-memory reads stay in program order: all of them first)
#pragma unroll
                for (int j = 0; j < NJ; ++j) v[i][j] = lds_f32(tp + (uint32_t)(i * S::EPI_PITCH + 32 * j) * 4u) + bj[j];
#pragma unroll
            for (int pi = 0; pi < CK - 1; ++pi) {
                float w[RB][NJ];
#pragma unroll
                for (int i = 0; i < RB; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) w[i][j] = lds_cluster_f32(peer[pi] + (uint32_t)(i * S::EPI_PITCH + 32 * j) * 4u);
#pragma unroll
                for (int i = 0; i < RB; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) v[i][j] += w[i][j];
                peer[pi] += RB * S::EPI_PITCH * 4;
            }
            if (act == ACT_ELU) {
#pragma unroll
                for (int i = 0; i < RB; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) v[i][j] = elu_fast(v[i][j]);
            } else if (act == ACT_RELU) {
#pragma unroll
                for (int i = 0; i < RB; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) v[i][j] = fmaxf(v[i][j], 0.f);
            } else if (act == ACT_TANH) {
#pragma unroll
                for (int i = 0; i < RB; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) v[i][j] = tanhf(v[i][j]);
            }
            if (has_d) {
#pragma unroll
                for (int i = 0; i < RB; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) { const float y = d[i][j]; v[i][j] *= y > thr ? 1.f : fmaf(fmaf(a2, y, a1), y, a0); }
            }
            if (is_add) {
#pragma unroll
                for (int i = 0; i < RB; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) v[i][j] += old[i][j];
            }
            float* cpi = cp;
#pragma unroll
            for (int i = 0; i < RB; ++i) {
#pragma unroll
                for (int j = 0; j < NJ; ++j)
                    if (cv[j] && (FULL || i < nrow)) { if (is_atomic) atomicAdd(cpi + 32 * j, v[i][j]); else cpi[32 * j] = v[i][j]; }
                cpi += g.ldc;
            }
            dp += (size_t)RB * g.ldd; cp += (size_t)RB * g.ldc; tp += RB * S::EPI_PITCH * 4;
        };
#pragma unroll 1
        for (int r0 = rbeg; r0 < rend; r0 += RB) {
            if (rend - r0 >= RB) batch(std::true_type{}, RB);
            else batch(std::false_type{}, rend - r0);
            if (r0 == rbeg) TC5_MARK(7);
        }
    }
    TC5_MARK(5);
    if (CK > 1) cluster_sync_all();                  // a peer may still be reading this CTA's tile
}

template <bool TA, bool TB, int TN, int CK>
inline cudaError_t launch_one(const GemmArgs& g, dim3 grid, cudaStream_t s) {
    constexpr int NS = 2;
    auto* kern = gemm_tc5_kernel<TA, TB, TN, NS, CK>;
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem5<TN, NS>::BYTES);
        configured = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(THREADS5); cfg.dynamicSmemBytes = Smem5<TN, NS>::BYTES; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = CK;
    cfg.attrs = at; cfg.numAttrs = CK > 1 ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, g);
}

// tn = 64 / 80; ck = 1 (grid.z = k-chunks of g.kchunk) or 2 / 4 (grid.z = ck, one cluster per tile; store / add outputs only)
template <bool TA, bool TB>
inline cudaError_t launch(const GemmArgs& g, int tn, int ck, dim3 grid, cudaStream_t s) {
    if (tn == 64) {
        if (ck == 1) return launch_one<TA, TB, 64, 1>(g, grid, s);
        if (ck == 2) return launch_one<TA, TB, 64, 2>(g, grid, s);
        return launch_one<TA, TB, 64, 4>(g, grid, s);
    }
    if (ck == 1) return launch_one<TA, TB, 80, 1>(g, grid, s);
    if (ck == 2) return launch_one<TA, TB, 80, 2>(g, grid, s);
    return launch_one<TA, TB, 80, 4>(g, grid, s);
}

}  // namespace tc5
