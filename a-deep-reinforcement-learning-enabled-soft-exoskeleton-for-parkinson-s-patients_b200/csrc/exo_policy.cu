// exo_policy.cu — fused policy forward pass on the 5th-generation tensor cores (include/exo_policy.h).
//
//   encoder.zs + actor of Agent/TD7_multi_agent_Pink_noise.py:66-105, evaluated as select_action does (:208-228), for
//   observations resident in HBM. ~1.0 MFLOP per observation in seven small GEMMs: the one GEMM-shaped piece of the
//   env -> agent loop, so this is where tcgen05 is used (the env step itself is HBM-bound 7x7 work).
//
// One persistent CTA per SM, 128 observations (= the 128 TMEM lanes = UMMA M) per tile, 320 threads:
//   warp 0 / lane 0   weight producer: cp.async.bulk (TMA engine, 1-D) global -> shared stages, mbarrier complete_tx
//   warp 1            TMEM allocation; one elected lane issues every tcgen05.mma (kind::tf32, cta_group::1, M=128,
//                     N = 160 + 144 per k-step) - the warp stays converged so the issue path is the uniform datapath
//   warps 2..9        epilogue: two warps per TMEM lane quarter (columns 0..159 / 160..303): tcgen05.ld 32 accumulator
//                     columns at a time, activation, round to TF32, store as the next layer's A operand in shared memory
// The biases ride in the GEMMs: every A operand carries a constant-1 column (index 300 of a hidden vector, 80 of an
// observation) and the packed weights hold the bias there, so the epilogues load nothing but the accumulator.
// Shared memory (226 KB): X = 128 x 320 activations (160 KB = ten 32-column k-atoms; the observations sit in its first
// three until the first hidden layer overwrites them), 3 weight stages of 20 KB (one k-atom of 160 or 144 weight rows =
// 4 MMAs), mbarriers. Operands use the K-major SWIZZLE_128B canonical layout (8 rows x 128 B atoms, SBO 1024), which the
// host packs the weights into once, so a stage is ONE contiguous bulk copy and needs no tensor map.
// Tensor memory: accumulator P = columns 0..303 of the CTA's 512.
//
// Per tile (G = MMAs, E = epilogue; hidden width 300 padded to 304 with zero weights/bias, so pads stay exactly 0):
//   E0 obs -> X | G1 X*l0' -> E1 a0 = AvgL1Norm -> L2 scratch | G2 X*zs1' -> E2 ELU -> X | G3 X*zs2' -> E3 ELU -> X |
//   G4 X*zs3' -> E4 zs = AvgL1Norm -> X | G5a P = X*l1[:,300:600]' | bulk-reload X <- a0, k-atom by k-atom behind G5a | G5b P += X*l1[:,0:300]' ->
//   E5 ReLU -> X | G6 X*l2' -> E6 ReLU -> X | G7 X*l3' (N=16) -> E7 tanh, clip, * max_action -> actions
// (a0 takes the detour through an L2-resident per-CTA scratch because X + a second 160 KB operand do not fit in
// shared memory and two 304-column accumulators do not fit in TMEM.)
// Tried and measured slower: a second low-half accumulator in the spare TMEM columns so that the low
// half of GEMM k+1 runs under epilogue k (all low-half MMAs of a layer then form ONE dependent accumulation chain:
// 0.250 ms instead of 0.227), and four 80/64-column chains per stage (per-MMA cost dominates: 0.284 ms).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "../../include/exo_env.h"
#include "../../include/exo_policy.h"
#include "exo_policy_fp32.h"

extern "C" const char* exo_last_error(void);
int exo_set_error(int code, const char* fmt, ...);     // exo_api.cu

#ifndef EXO_POLICY_PAIR_DEFAULT
#define EXO_POLICY_PAIR_DEFAULT false
#endif

#define PCK(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t _e = (call);                                                                           \
        if (_e != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
    } while (0)


namespace {

constexpr int SD = EXO_POLICY_STATE_DIM, HD = EXO_POLICY_HIDDEN_DIM, AD = EXO_POLICY_ACTION_DIM;
constexpr int HP = 304;                     // hidden width padded: multiple of 8 (MMA K of tf32) and of 16 (MMA N)
constexpr int SP = 88;                      // observation width padded to hold the constant-1 column (index 80)
constexpr int OUTP = 16;                    // l3 rows padded to the smallest MMA N
constexpr int TILE_M = 128;
constexpr int KS_S = SP / 8;                // 11 k-steps over the observation
constexpr int KS_H = HP / 8;                // 38 k-steps over a hidden vector
constexpr int N_LO = 160, N_HI = HP - N_LO; // the two MMAs of a k-step (N <= 256 per instruction) = the two epilogue halves
// Operands are K-major SWIZZLE_128B: a swizzle atom is 8 rows x 128 B (32 tf32 of K = 4 k-steps), atoms of one K range
// stacked along the rows (SBO 1024), K ranges ("k-atoms") one after the other.
constexpr int ATOMS_S = (SP + 31) / 32;     // 3 k-atoms over the observation (the last one 3 k-steps)
constexpr int ATOMS_H = (HP + 31) / 32;     // 10 k-atoms over a hidden vector (the last one 2 k-steps)
constexpr int A_ATOM_BYTES = TILE_M * 128;  // 16 384: one k-atom of the 128-row activation tile
constexpr int X_BYTES = ATOMS_H * A_ATOM_BYTES;              // 163 840
constexpr int B_LO_BYTES = N_LO * 128, B_HI_BYTES = N_HI * 128;   // 20 480 / 18 432: one k-atom of weight rows 0..159 / 160..303
constexpr int SLOT_BYTES = B_LO_BYTES;      // a weight stage = one (k-atom, row half); all of l3 (10 k-atoms x 16 rows) is one stage too
static_assert(ATOMS_H * OUTP * 128 == SLOT_BYTES, "l3 fills exactly one stage");
#ifndef EXO_POLICY_NSLOT
#define EXO_POLICY_NSLOT 3
#endif
constexpr int NSLOT = EXO_POLICY_NSLOT;   // weight stages in flight (tuning builds: -DEXO_POLICY_NSLOT=n)
constexpr int N_PHASE = 8;
constexpr int PACK_BYTES = (2 * ATOMS_S + 5 * ATOMS_H) * (B_LO_BYTES + B_HI_BYTES) + SLOT_BYTES;     // 2 199 552

constexpr int OFF_X = 0;
constexpr int OFF_W = OFF_X + X_BYTES;
constexpr int OFF_BAR = OFF_W + NSLOT * SLOT_BYTES;
constexpr int N_BAR = 3 * NSLOT + 4 + ATOMS_H;                     // full[], empty[], acc_full, a_ready, xr_full, xa_free[]; pair mode: pfull[], pxr
// CTA-pair mode (cta_group::2, M = 256 = two 128-row tiles): each CTA of the pair stages HALF of the weight rows of an MMA
// (rows 0..79 | 80..159 of the low half, 160..231 | 232..303 of the high half); one stage = one k-atom of both = 8 MMAs.
constexpr int B2_LO_BYTES = (N_LO / 2) * 128, B2_HI_BYTES = (N_HI / 2) * 128;     // 10 240 / 9 216
constexpr int B2_STAGE_BYTES = B2_LO_BYTES + B2_HI_BYTES;                         // 19 456 per CTA and k-atom
constexpr int B2_L3_BYTES = ATOMS_H * 8 * 128;                                    // 10 240: l3, 8 of its 16 (padded) rows per CTA
static_assert(B2_STAGE_BYTES <= SLOT_BYTES && B2_LO_BYTES % 1024 == 0, "pair-mode stage layout");
constexpr int OFF_PART = OFF_BAR + 8 * N_BAR + 16;                 // AvgL1Norm partial sums [2][128] floats
constexpr int SMEM_BYTES = OFF_PART + 2 * TILE_M * 4;
static_assert(SMEM_BYTES <= 232448, "over the 227 KB per-CTA shared memory limit");
static_assert(N_LO % 32 == 0 && N_HI % 16 == 0 && N_HI > 128, "epilogue column split");

constexpr int N_EPI = 256;
constexpr int N_THREADS = 64 + N_EPI;
constexpr int OBS_V4_PER_THREAD = (TILE_M * SD / 4) / N_EPI;       // 10

// ---- PTX wrappers ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
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
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (!done && (spins & 1023u) == 1023u) {
            const long long t = clock64();
            if (t0 == 0) t0 = t;
            else if (t - t0 > 4000000000ll) __trap();
        }
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// ---- CTA-pair (cluster of 2) helpers
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t cta) {
    asm volatile("{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\t"
                 "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(bar), "r"(cta) : "memory");
}
// wait on a barrier that the peer CTA arrives on (cluster-scope acquire)
__device__ __forceinline__ void mbar_wait_cl(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (uint32_t spins = 0; !done; ++spins) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (!done && (spins & 1023u) == 1023u) {
            const long long t = clock64();
            if (t0 == 0) t0 = t;
            else if (t - t0 > 4000000000ll) __trap();
        }
    }
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {      // arrives on `bar` in BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((unsigned short)3)
                 : "memory");
}
// M = 256 over the CTA pair: each CTA multiplies its own 128 rows of A by B = [this CTA's N/2 rows ; the peer's N/2 rows]
__device__ __forceinline__ void tc_mma_tf32_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, both operands K-major, tf32 inputs, fp32 accumulate
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]),
          "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]),
          "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// cvt.rna.tf32.f32 (round to nearest, ties away from zero, 10-bit mantissa) as two integer operations
__device__ __forceinline__ float to_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u); }
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, %1;\n\tselp.b32 %0, 1, 0, px;\n\t}" : "=r"(pred) : "r"(0xFFFFFFFFu));
    return pred != 0;
}

// Shared-memory matrix descriptor (sm_100 UMMA): start>>4 [0,14), LBO>>4 [16,30) (unused by a K-major swizzled operand
// whose MMA K extent lies inside one swizzle atom; set to 1), SBO>>4 [32,46) = distance of 8-row groups = 1024,
// version 1 [46,48), layout type [61,64) = 2: SWIZZLE_128B. A k-step inside the atom is selected by adding its 32 B to
// the start address. (Measured, tools/ubench/mma_fetch.cu: 159 cycles per N=160+144 k-step with this layout, 201 with
// SWIZZLE_32B; the first version's no-swizzle layout was slower still.)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
    return (uint64_t)((addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Byte offset of 16-byte chunk `c` (4 tf32 of K; c = k / 4) of row `r` in an operand whose k-atoms are `atom_bytes` apart:
// rows 128 B apart inside an 8-row atom, atoms of later rows 1024 B further, and address bits [4,7) XORed with bits
// [7,10), i.e. chunk (c & 7) of row r sits at position (c & 7) ^ (r & 7).
__host__ __device__ __forceinline__ uint32_t op_off(uint32_t r, uint32_t c, uint32_t atom_bytes) {
    return (c >> 3) * atom_bytes + (r >> 3) * 1024u + (r & 7u) * 128u + (((c & 7u) ^ (r & 7u)) << 4);
}
// Instruction descriptor: D fp32 (1<<4), A and B tf32 (2<<7, 2<<10), both K-major, N>>3 at bit 17, M>>4 at bit 24.
__host__ __device__ constexpr uint32_t idesc_tf32(int n, int m = TILE_M) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

struct PolicyArgs {
    const float* obs;            // [n, 80]
    const unsigned char* pack;   // the weight stages of one tile, in consumption order (PACK_BYTES)
    unsigned char* scratch;      // gridDim.x * X_BYTES (a0 in operand layout)
    float* actions;              // [n, 7]
    float* zs;                   // [n, 300] or null
    long long n;
    float max_action;
    int n_tiles;
    long long* dbg;              // exo_policy_profile: clock64 stamps of CTA 0's first tile, or null
};

// k-atoms per phase (G1, G2, G3, G4, G5a, G5b, G6, G7) and k-steps in the last one
__device__ __forceinline__ int ph_atoms(int ph) { return ph < 2 ? ATOMS_S : ATOMS_H; }
__device__ __forceinline__ int ph_last_ks(int ph) { return ph < 2 ? KS_S - 4 * (ATOMS_S - 1) : KS_H - 4 * (ATOMS_H - 1); }

// ---- epilogues: a thread owns row r of the tile and the column half `half` (0: 0..159, 1: 160..303) ---------------
// Stores 4 consecutive columns (chunk index c of the row) of an operand-layout row.
__device__ __forceinline__ void st_chunk(unsigned char* row, uint32_t rsw, int c, float a, float b, float cc, float d) {
    *reinterpret_cast<float4*>(row + (c >> 3) * A_ATOM_BYTES + ((((uint32_t)c & 7u) ^ rsw) << 4)) = make_float4(a, b, cc, d);
}

// ACT 0: ELU, 1: ReLU, 2: multiply by `scale` (second pass of AvgL1Norm). Processes NCOL (32 or 16) accumulator columns
// starting at column col0 and writes them, rounded to TF32, into the operand row; column 300 becomes the constant 1.
template <int ACT, int NCOL>
__device__ __forceinline__ void epi_block(const uint32_t* v, int col0, unsigned char* row, uint32_t rsw, float scale, float* zs_row) {
#pragma unroll
    for (int j4 = 0; j4 < NCOL / 4; ++j4) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float x = __uint_as_float(v[j4 * 4 + j]);
            if (ACT == 0) { const float e = __expf(fminf(x, 0.f)) - 1.f; x = x > 0.f ? x : e; }
            else if (ACT == 1) x = fmaxf(x, 0.f);
            else x *= scale;
            o[j] = x;
        }
        if (ACT == 2 && zs_row) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (col0 + j4 * 4 + j < HD) zs_row[col0 + j4 * 4 + j] = o[j];
        }
        if (col0 + j4 * 4 == HD) o[0] = 1.0f;              // the bias column of the next GEMM
        st_chunk(row, rsw, (col0 >> 2) + j4, to_tf32(o[0]), to_tf32(o[1]), to_tf32(o[2]), to_tf32(o[3]));
    }
}

template <int ACT>
__device__ __forceinline__ void epi_half(uint32_t taddr, int half, unsigned char* row, uint32_t rsw, float scale, float* zs_row) {
    uint32_t v[32];
    const int c0 = half * N_LO;
#pragma unroll 1
    for (int b = 0; b < 4; ++b) {
        tmem_ld32(taddr + c0 + b * 32, v);
        tmem_ld_wait();
        epi_block<ACT, 32>(v, c0 + b * 32, row, rsw, scale, zs_row);
    }
    if (half == 0) {
        tmem_ld32(taddr + 128, v);
        tmem_ld_wait();
        epi_block<ACT, 32>(v, 128, row, rsw, scale, zs_row);
    } else {
        tmem_ld16(taddr + N_LO + 128, v);
        tmem_ld_wait();
        epi_block<ACT, 16>(v, N_LO + 128, row, rsw, scale, zs_row);
    }
}

// sum of |accumulator| over this thread's column half
__device__ __forceinline__ float epi_abs_sum(uint32_t taddr, int half) {
    uint32_t v[32];
    const int c0 = half * N_LO;
    float s = 0.f;
#pragma unroll 1
    for (int b = 0; b < 4; ++b) {
        tmem_ld32(taddr + c0 + b * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) s += fabsf(__uint_as_float(v[j]));
    }
    if (half == 0) {
        tmem_ld32(taddr + 128, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) s += fabsf(__uint_as_float(v[j]));
    } else {
        tmem_ld16(taddr + N_LO + 128, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) s += fabsf(__uint_as_float(v[j]));
    }
    return s;
}

// AvgL1Norm (:59-60): x / clamp(mean|x|, 1e-8); the two column halves of a row exchange their partial sums through
// shared memory (pad columns are exactly 0 and do not touch the sum). Returns 1 / denominator.
__device__ __forceinline__ float avgl1_scale(uint32_t taddr, int half, int r, float* part) {
    const float s = epi_abs_sum(taddr, half);
    part[half * TILE_M + r] = s;
    asm volatile("bar.sync 2, 256;" ::: "memory");
    const float tot = s + part[(half ^ 1) * TILE_M + r];
    asm volatile("bar.sync 2, 256;" ::: "memory");       // part[] may be rewritten by the next normalisation
    return 1.0f / fmaxf(tot * (1.0f / (float)HD), 1e-8f);
}

#define STAMP(i) do { if (a.dbg && tile == 0 && et == 0) a.dbg[i] = clock64(); } while (0)

// PAIR = false: every CTA works alone (cta_group::1). PAIR = true: launched in clusters of 2; CTA rank 0 of a pair issues
// tcgen05.mma.cta_group::2 for both 128-row tiles (M = 256), each CTA streams half of the weight rows, rank 1's warp 1
// relays "my half has landed" to rank 0. Epilogues, operand X, scratch and the tile <-> CTA mapping are the same.
template <bool PAIR>
__global__ void __launch_bounds__(N_THREADS, 1) exo_policy_kernel(const PolicyArgs a) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = PAIR ? cluster_rank() : 0u;
    const int n_iter = (a.n_tiles + (int)gridDim.x - 1) / (int)gridDim.x;     // the same for every CTA: a pair stays in step
    const uint32_t sbase = smem_u32(smem);
    // barriers: full[NSLOT], empty[NSLOT], acc_full, a_ready, xr_full, xa_free[ATOMS_H]; then the TMEM base address
    const uint32_t bar_full = sbase + OFF_BAR, bar_empty = bar_full + 8 * NSLOT;
    const uint32_t bar_acc = bar_empty + 8 * NSLOT, bar_aready = bar_acc + 8, bar_xr = bar_aready + 8, bar_xafree = bar_xr + 8;
    const uint32_t bar_pfull = bar_xafree + 8 * ATOMS_H, bar_pxr = bar_pfull + 8 * NSLOT;   // pair mode, used in rank 0: the peer's full[] / xr
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + OFF_BAR + 8 * N_BAR);

    if (tid == 0) {
        for (int s = 0; s < NSLOT; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        mbar_init(bar_acc, 1);
        mbar_init(bar_aready, PAIR ? 2 * N_EPI : N_EPI);          // pair mode: both CTAs' epilogue threads arrive at rank 0
        mbar_init(bar_xr, 1);
        for (int g = 0; g < ATOMS_H; ++g) mbar_init(bar_xafree + 8 * g, 1);
        for (int s = 0; s < NSLOT; ++s) mbar_init(bar_pfull + 8 * s, 1);
        mbar_init(bar_pxr, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync();          // both CTAs' barriers exist before anything arrives on them remotely
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===== weight producer ==========================================================================
        if (lane == 0) {
            uint32_t slot = 0, par = 1;
            auto stage = [&](const unsigned char* src, uint32_t bytes) {
                mbar_wait(bar_empty + 8 * slot, par);
#ifdef EXO_POLICY_TINYLOAD       // timing experiment only (wrong results): is the k-step paced by the bytes a stage brings in?
                bytes = 16;
#endif
                mbar_expect_tx(bar_full + 8 * slot, bytes);
                bulk_g2s(sbase + OFF_W + slot * SLOT_BYTES, src, bytes, bar_full + 8 * slot);
                if (++slot == NSLOT) { slot = 0; par ^= 1; }
            };
            for (int it = 0; it < n_iter; ++it) {
                const unsigned char* src = a.pack;
#pragma unroll 1
                for (int ph = 0; ph < 7; ++ph) {
                    const int na = ph_atoms(ph);
#pragma unroll 1
                    for (int at = 0; at < na; ++at) {
                        if (PAIR) {                              // this CTA's half of the k-atom: [rank 0 block][rank 1 block]
                            stage(src + rank * B2_STAGE_BYTES, B2_STAGE_BYTES); src += 2 * B2_STAGE_BYTES;
                        } else {
                            stage(src, B_LO_BYTES); src += B_LO_BYTES;
                            stage(src, B_HI_BYTES); src += B_HI_BYTES;
                        }
                    }
                }
                if (PAIR) stage(src + rank * B2_L3_BYTES, B2_L3_BYTES);
                else stage(src, SLOT_BYTES);
            }
        }
    } else if (warp == 1 && PAIR) {
        uint32_t slot = 0, par = 0, par_ar = 0, par_xr = 0;
        if (rank == 0) {
            // ===== pair mode, rank 0: MMA issuer for both tiles =============================================
            constexpr uint32_t IDLO = idesc_tf32(N_LO, 256), IDHI = idesc_tf32(N_HI, 256), ID16 = idesc_tf32(OUTP, 256);
            const uint64_t a_desc0 = smem_desc(sbase + OFF_X);
            const uint64_t b_desc0 = smem_desc(sbase + OFF_W);
            for (int it = 0; it < n_iter; ++it) {
#pragma unroll 1
                for (int ph = 0; ph < N_PHASE; ++ph) {
                    if (ph == 5) { mbar_wait(bar_xr, par_xr); mbar_wait_cl(bar_pxr, par_xr); par_xr ^= 1; }
                    else { mbar_wait_cl(bar_aready, par_ar); par_ar ^= 1; }
                    tc_fence_after();
                    const bool stamp = a.dbg && it == 0 && blockIdx.x == 0 && lane == 0;
                    if (stamp) a.dbg[2 * ph] = clock64();
                    const int na = ph_atoms(ph), last_ks = ph_last_ks(ph);
                    if (ph < 7) {
#pragma unroll 1
                        for (int at = 0; at < na; ++at) {          // stage = one k-atom: both column halves, 8 MMAs
                            mbar_wait(bar_full + 8 * slot, par);
                            mbar_wait_cl(bar_pfull + 8 * slot, par);
                            tc_fence_after();
                            const uint64_t ad = a_desc0 + (uint64_t)(at * (A_ATOM_BYTES >> 4));
                            const uint64_t bd = b_desc0 + (uint64_t)(slot * (SLOT_BYTES >> 4));
                            const int nk = at == na - 1 ? last_ks : 4;
                            if (elect_one()) {
                                const uint32_t acc0 = (at > 0 || ph == 5) ? 1u : 0u;
                                // k-step by k-step, the two column halves alternating: consecutive MMAs never share an accumulator
                                // (tools/ubench/mma_fetch.cu: 159 cycles per alternating pair, ~200 per MMA in a one-accumulator chain)
#pragma unroll
                                for (int j = 0; j < 4; ++j)
                                    if (j < nk) {
                                        tc_mma_tf32_pair(tmem, ad + (uint64_t)(2 * j), bd + (uint64_t)(2 * j), IDLO, (j > 0) ? 1u : acc0);
                                        tc_mma_tf32_pair(tmem + N_LO, ad + (uint64_t)(2 * j), bd + (uint64_t)((B2_LO_BYTES >> 4) + 2 * j), IDHI,
                                                         (j > 0) ? 1u : acc0);
                                    }
                                tc_commit_pair(bar_empty + 8 * slot);          // both CTAs' stages are free once these MMAs have read them
                                if (ph == 4) tc_commit_pair(bar_xafree + 8 * at);   // G5a is done with this k-atom of X (in both CTAs)
                            }
                            __syncwarp();
                            if (++slot == NSLOT) { slot = 0; par ^= 1; }
                        }
                    } else {                                        // l3: 8 of its 16 padded rows from each CTA, all k-atoms in one stage
                        mbar_wait(bar_full + 8 * slot, par);
                        mbar_wait_cl(bar_pfull + 8 * slot, par);
                        tc_fence_after();
                        const uint64_t bd = b_desc0 + (uint64_t)(slot * (SLOT_BYTES >> 4));
                        if (elect_one()) {
#pragma unroll 1
                            for (int at = 0; at < ATOMS_H; ++at) {
                                const int nk = at == ATOMS_H - 1 ? last_ks : 4;
#pragma unroll
                                for (int j = 0; j < 4; ++j)
                                    if (j < nk)
                                        tc_mma_tf32_pair(tmem, a_desc0 + (uint64_t)(at * (A_ATOM_BYTES >> 4) + 2 * j),
                                                         bd + (uint64_t)(at * (8 * 128 >> 4) + 2 * j), ID16, (at > 0 || j > 0) ? 1u : 0u);
                            }
                            tc_commit_pair(bar_empty + 8 * slot);
                        }
                        __syncwarp();
                        if (++slot == NSLOT) { slot = 0; par ^= 1; }
                    }
                    if (ph != 4) {
                        if (elect_one()) tc_commit_pair(bar_acc);                // both CTAs' epilogues may read their accumulators
                        __syncwarp();
                    }
                    if (stamp) a.dbg[2 * ph + 1] = clock64();
                }
            }
        } else {
            // ===== pair mode, rank 1: tell rank 0 when this CTA's weight halves / reloaded a0 have landed ===========
            for (int it = 0; it < n_iter; ++it) {
#pragma unroll 1
                for (int ph = 0; ph < N_PHASE; ++ph) {
                    if (ph == 5) {
                        mbar_wait(bar_xr, par_xr); par_xr ^= 1;
                        if (lane == 0) mbar_arrive_remote(bar_pxr, 0);
                        __syncwarp();
                    }
                    const int nst = ph < 7 ? ph_atoms(ph) : 1;
#pragma unroll 1
                    for (int st = 0; st < nst; ++st) {
                        mbar_wait(bar_full + 8 * slot, par);
                        if (lane == 0) mbar_arrive_remote(bar_pfull + 8 * slot, 0);
                        __syncwarp();
                        if (++slot == NSLOT) { slot = 0; par ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the whole warp walks the pipeline, one elected lane issues =====================
        uint32_t slot = 0, par = 0, par_ar = 0, par_xr = 0;
        constexpr uint32_t IDLO = idesc_tf32(N_LO), IDHI = idesc_tf32(N_HI), ID16 = idesc_tf32(OUTP);
        const uint64_t a_desc0 = smem_desc(sbase + OFF_X);
        const uint64_t b_desc0 = smem_desc(sbase + OFF_W);
        for (int it = 0; it < n_iter; ++it) {
            const int tile = blockIdx.x + it * gridDim.x;
#pragma unroll 1
            for (int ph = 0; ph < N_PHASE; ++ph) {
                if (ph == 5) { mbar_wait(bar_xr, par_xr); par_xr ^= 1; }
                else { mbar_wait(bar_aready, par_ar); par_ar ^= 1; }
                tc_fence_after();
                const bool stamp = a.dbg && tile == 0 && lane == 0;
                if (stamp) a.dbg[2 * ph] = clock64();
                const int na = ph_atoms(ph), last_ks = ph_last_ks(ph);
                if (ph < 7) {
#pragma unroll 1
                    for (int st = 0; st < 2 * na; ++st) {          // stage = (k-atom st/2, weight-row half st&1)
                        mbar_wait(bar_full + 8 * slot, par);
                        tc_fence_after();
                        const int at = st >> 1, hi = st & 1;
                        const uint64_t ad = a_desc0 + (uint64_t)(at * (A_ATOM_BYTES >> 4));
                        const uint64_t bd = b_desc0 + (uint64_t)(slot * (SLOT_BYTES >> 4));
                        const int nk = at == na - 1 ? last_ks : 4;
                        if (elect_one()) {
                            const uint32_t d = tmem + (hi ? N_LO : 0);
                            const uint32_t id = hi ? IDHI : IDLO;
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (j < nk) tc_mma_tf32(d, ad + (uint64_t)(2 * j), bd + (uint64_t)(2 * j), id, (at > 0 || j > 0 || ph == 5) ? 1u : 0u);
                            tc_commit(bar_empty + 8 * slot);      // frees the stage once these MMAs have read it
                            if (ph == 4 && hi) tc_commit(bar_xafree + 8 * at);    // G5a is done with this k-atom of X: a0 may come back
                        }
                        __syncwarp();
                        if (++slot == NSLOT) { slot = 0; par ^= 1; }
                    }
                } else {                                            // l3: all ten k-atoms of its 16 rows in one stage
                    mbar_wait(bar_full + 8 * slot, par);
                    tc_fence_after();
                    const uint64_t bd = b_desc0 + (uint64_t)(slot * (SLOT_BYTES >> 4));
                    if (elect_one()) {
#pragma unroll 1
                        for (int at = 0; at < ATOMS_H; ++at) {
                            const int nk = at == ATOMS_H - 1 ? last_ks : 4;
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (j < nk)
                                    tc_mma_tf32(tmem, a_desc0 + (uint64_t)(at * (A_ATOM_BYTES >> 4) + 2 * j),
                                                bd + (uint64_t)(at * (OUTP * 128 >> 4) + 2 * j), ID16, (at > 0 || j > 0) ? 1u : 0u);
                        }
                        tc_commit(bar_empty + 8 * slot);
                    }
                    __syncwarp();
                    if (++slot == NSLOT) { slot = 0; par ^= 1; }
                }
                if (ph != 4) {                                  // (G5a's accumulator is completed by G5b)
                    if (elect_one()) tc_commit(bar_acc);
                    __syncwarp();
                }
                if (stamp) a.dbg[2 * ph + 1] = clock64();
            }
        }
    } else {
        // ===== epilogue warps ===========================================================================
        const int et = tid - 64;                       // 0..255
        const int half = et >> 7;                      // column half
        const int q = warp & 3;                        // TMEM lane quarter this warp may read
        const int r = q * 32 + lane;                   // accumulator row = observation within the tile
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
        const uint32_t rsw = (uint32_t)r & 7u;         // chunk swizzle of this row
        unsigned char* xrow = smem + OFF_X + (r >> 3) * 1024 + (r & 7) * 128;
        unsigned char* scratch = a.scratch + (size_t)blockIdx.x * X_BYTES;
        unsigned char* srow = scratch + (r >> 3) * 1024 + (r & 7) * 128;
        float* part = reinterpret_cast<float*>(smem + OFF_PART);
        uint32_t par_acc = 0, par_xf = 0;
        float4 ob[OBS_V4_PER_THREAD];                  // this thread's share of the next tile's observations
        auto load_obs = [&](int tile) {
            const long long row0 = (long long)tile * TILE_M;
#pragma unroll
            for (int i = 0; i < OBS_V4_PER_THREAD; ++i) {
                const int idx = i * N_EPI + et, rr = idx / (SD / 4), c = idx % (SD / 4);
                ob[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (tile < a.n_tiles && row0 + rr < a.n) ob[i] = __ldg(reinterpret_cast<const float4*>(a.obs + (row0 + rr) * SD) + c);
            }
        };
        // "the previous epilogue is done" goes to the MMA issuer: this CTA's, or in pair mode the one in rank 0
        auto arrive_aready = [&]() {
            if (PAIR && rank != 0) mbar_arrive_remote(bar_aready, 0);
            else mbar_arrive(bar_aready);
        };
        load_obs(blockIdx.x);
        for (int it = 0; it < n_iter; ++it) {
            const int tile = blockIdx.x + it * gridDim.x;          // may lie past the end in the last round: every row is masked then
            const long long row0 = (long long)tile * TILE_M;
            STAMP(16);
            // E0: observations -> the first 22 chunks of X's rows (80 values, the constant 1, 7 zeros), rounded to TF32; both
            // layers that read them (l0, zs1) are done before E2 overwrites X. Then prefetch the next tile's.
#pragma unroll
            for (int i = 0; i < OBS_V4_PER_THREAD; ++i) {
                const int idx = i * N_EPI + et, rr = idx / (SD / 4), c = idx % (SD / 4);
                *reinterpret_cast<float4*>(smem + OFF_X + op_off(rr, c, A_ATOM_BYTES)) =
                    make_float4(to_tf32(ob[i].x), to_tf32(ob[i].y), to_tf32(ob[i].z), to_tf32(ob[i].w));
            }
            {
                const int rr = et & 127, c = SD / 4 + half;           // chunk 20: (1,0,0,0); chunk 21: zeros
                *reinterpret_cast<float4*>(smem + OFF_X + op_off(rr, c, A_ATOM_BYTES)) = make_float4(half ? 0.f : 1.f, 0.f, 0.f, 0.f);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            STAMP(17);
            arrive_aready();
            load_obs(tile + gridDim.x);
            // E1: a0 = AvgL1Norm(l0(obs)) -> scratch
            mbar_wait(bar_acc, par_acc); par_acc ^= 1; tc_fence_after();
            STAMP(18);
            {
                const float sc = avgl1_scale(taddr, half, r, part);
                epi_half<2>(taddr, half, srow, rsw, sc, nullptr);
            }
            tc_fence_before();
            STAMP(19);
            arrive_aready();                         // the accumulator is drained: zs1 may start while the stores land
            __threadfence();
            fence_proxy_async_all();
            asm volatile("bar.sync 1, 256;" ::: "memory");   // all of a0 is out (and visible to the bulk copies) before the reload
            // E2, E3: ELU
            mbar_wait(bar_acc, par_acc); par_acc ^= 1; tc_fence_after();
            STAMP(20);
            epi_half<0>(taddr, half, xrow, rsw, 0.f, nullptr);
            fence_proxy_async_smem(); tc_fence_before();
            STAMP(21);
            arrive_aready();
            mbar_wait(bar_acc, par_acc); par_acc ^= 1; tc_fence_after();
            STAMP(22);
            epi_half<0>(taddr, half, xrow, rsw, 0.f, nullptr);
            fence_proxy_async_smem(); tc_fence_before();
            STAMP(23);
            arrive_aready();
            // E4: zs = AvgL1Norm(zs3(.))
            mbar_wait(bar_acc, par_acc); par_acc ^= 1; tc_fence_after();
            STAMP(24);
            {
                const float sc = avgl1_scale(taddr, half, r, part);
                epi_half<2>(taddr, half, xrow, rsw, sc, (a.zs && row0 + r < a.n) ? a.zs + (row0 + r) * HD : nullptr);
            }
            fence_proxy_async_smem(); tc_fence_before();
            STAMP(25);
            arrive_aready();
            // as G5a finishes reading each k-atom of zs out of X, bring that k-atom of a0 back (one bulk copy each)
            if (et == 0) {
                mbar_expect_tx(bar_xr, X_BYTES);
                for (int g = 0; g < ATOMS_H; ++g) {
                    mbar_wait(bar_xafree + 8 * g, par_xf);
                    bulk_g2s(sbase + OFF_X + g * A_ATOM_BYTES, scratch + (size_t)g * A_ATOM_BYTES, A_ATOM_BYTES, bar_xr);
                }
            }
            par_xf ^= 1;
            // E5, E6: ReLU
            mbar_wait(bar_acc, par_acc); par_acc ^= 1; tc_fence_after();
            STAMP(26);
            epi_half<1>(taddr, half, xrow, rsw, 0.f, nullptr);
            fence_proxy_async_smem(); tc_fence_before();
            STAMP(27);
            arrive_aready();
            mbar_wait(bar_acc, par_acc); par_acc ^= 1; tc_fence_after();
            STAMP(28);
            epi_half<1>(taddr, half, xrow, rsw, 0.f, nullptr);
            fence_proxy_async_smem(); tc_fence_before();
            STAMP(29);
            arrive_aready();
            // E7: actions (column half 0 only; the other warps just keep the barrier phases)
            mbar_wait(bar_acc, par_acc); par_acc ^= 1; tc_fence_after();
            STAMP(30);
            if (half == 0) {
                uint32_t v[16];
                tmem_ld16(taddr, v);
                tmem_ld_wait();
                if (row0 + r < a.n) {
                    float* out = a.actions + (row0 + r) * AD;
#pragma unroll
                    for (int j = 0; j < AD; ++j) out[j] = fminf(fmaxf(tanhf(__uint_as_float(v[j])), -1.f), 1.f) * a.max_action;
                }
            }
            STAMP(31);
            tc_fence_before();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync();          // the peer may still be reading this CTA's weight stages / arriving on its barriers
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

// ---- host side ------------------------------------------------------------------------------------------
inline float round_tf32(float x) {            // cvt.rna.tf32.f32: round to nearest, ties away from zero, 10-bit mantissa
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7F800000u) != 0x7F800000u) u = (u + 0x1000u) & 0xFFFFE000u;
    memcpy(&x, &u, 4);
    return x;
}

// Packs W[n_rows, ld] (columns col0 .. col0+k_real-1; `bias`, if given, becomes column k_real - the A operand holds the
// constant 1 there) as the kernel consumes it: for each k-atom (32 columns) the rows 0..n_lo-1 as one K-major
// SWIZZLE_128B block, then the rows n_lo..n_lo+n_hi-1 (n_hi may be 0). Values rounded to TF32. Returns the bytes written.
size_t pack_matrix(unsigned char* dst, const float* W, const float* bias, int n_rows, int ld, int col0, int k_real, int n_lo, int n_hi,
                   int atoms) {
    const size_t atom_bytes = (size_t)(n_lo + n_hi) * 128;
    memset(dst, 0, atoms * atom_bytes);
    for (int n = 0; n < n_rows; ++n)
        for (int k = 0; k <= k_real; ++k) {
            if (k == k_real && !bias) break;
            const int hi = n >= n_lo, nl = hi ? n - n_lo : n;
            const size_t off = (size_t)(k / 32) * atom_bytes + (hi ? (size_t)n_lo * 128 : 0) + op_off((uint32_t)nl, (uint32_t)((k % 32) / 4), 0u) + (k % 4) * 4;
            const float v = round_tf32(k < k_real ? W[(size_t)n * ld + col0 + k] : bias[n]);
            memcpy(dst + off, &v, 4);
        }
    return atoms * atom_bytes;
}

// The same matrix for the CTA-pair kernel: per k-atom [rank 0 block][rank 1 block], a block = this CTA's half of the low
// rows (n_lo/2) followed by its half of the high rows (n_hi/2). With n_hi == 0 (l3) the blocks are per rank: all k-atoms
// of its n_lo/2 rows. Returns the bytes written.
size_t pack_matrix_pair(unsigned char* dst, const float* W, const float* bias, int n_rows, int ld, int col0, int k_real, int n_lo, int n_hi,
                        int atoms) {
    const int lo2 = n_lo / 2, hi2 = n_hi / 2;
    const size_t blk = (size_t)(lo2 + hi2) * 128;
    memset(dst, 0, 2 * atoms * blk);
    for (int n = 0; n < n_rows; ++n)
        for (int k = 0; k <= k_real; ++k) {
            if (k == k_real && !bias) break;
            const int hi = n >= n_lo, m = hi ? n - n_lo : n, half = hi ? hi2 : lo2;
            const int rank = m / half, nl = m % half, at = k / 32;
            const size_t base = n_hi ? ((size_t)at * 2 + rank) * blk + (hi ? (size_t)lo2 * 128 : 0)
                                     : (size_t)rank * atoms * blk + (size_t)at * blk;
            const size_t off = base + op_off((uint32_t)nl, (uint32_t)((k % 32) / 4), 0u) + (k % 4) * 4;
            const float v = round_tf32(k < k_real ? W[(size_t)n * ld + col0 + k] : bias[n]);
            memcpy(dst + off, &v, 4);
        }
    return 2 * atoms * blk;
}

}  // namespace

struct ExoPolicy {
    int device, n_sm;
    bool loaded;
    bool pair;                       // CTA-pair kernel (cta_group::2); EXO_POLICY_PAIR=0 selects the single-CTA kernel
    unsigned char *pack, *pack2, *scratch;
    ExoPolicyF32* f32;               // the float32 path (exo_policy_fp32.cu)
    int64_t launches;
    long long* dbg;
};

extern "C" int exo_policy_create(int32_t device, ExoPolicy** out) {
    if (!out) return exo_set_error(EXO_ERR_ARG, "exo_policy_create: null out");
    int ndev = 0;
    PCK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return exo_set_error(EXO_ERR_CUDA, "device %d not present: this library has no CPU path", device);
    cudaDeviceProp prop;
    PCK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return exo_set_error(EXO_ERR_CUDA, "device %d is sm_%d%d; libexo_env.so carries sm_100a code only", device, prop.major, prop.minor);
    PCK(cudaSetDevice(device));
    ExoPolicy* p = new ExoPolicy();
    p->device = device; p->n_sm = prop.multiProcessorCount; p->loaded = false; p->launches = 0;
    p->pack = p->pack2 = p->scratch = nullptr; p->dbg = nullptr; p->f32 = nullptr;
    const char* pm = getenv("EXO_POLICY_PAIR");
    p->pair = pm ? (pm[0] != '0') : EXO_POLICY_PAIR_DEFAULT;
    cudaError_t e = cudaMalloc((void**)&p->pack, PACK_BYTES);
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->pack2, PACK_BYTES);
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->scratch, (size_t)p->n_sm * X_BYTES);
    if (e == cudaSuccess) e = cudaMemset(p->scratch, 0, (size_t)p->n_sm * X_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(exo_policy_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(exo_policy_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) {
        exo_policy_destroy(p);
        return exo_set_error(EXO_ERR_CUDA, "exo_policy_create: %s", cudaGetErrorString(e));
    }
    if (int rc = exo_policy_f32_create(device, p->n_sm, &p->f32)) {
        exo_policy_destroy(p);
        return rc;
    }
    *out = p;
    return EXO_OK;
}

extern "C" int exo_policy_destroy(ExoPolicy* p) {
    if (!p) return EXO_OK;
    cudaSetDevice(p->device);
    cudaFree(p->pack); cudaFree(p->pack2); cudaFree(p->scratch);
    exo_policy_f32_destroy(p->f32);
    delete p;
    return EXO_OK;
}

extern "C" int exo_policy_set_weights(ExoPolicy* p, const ExoPolicyWeights* w, int32_t state_dim, int32_t hidden_dim, int32_t action_dim) {
    if (!p || !w) return exo_set_error(EXO_ERR_ARG, "exo_policy_set_weights: null argument");
    if (state_dim != SD || hidden_dim != HD || action_dim != AD)
        return exo_set_error(EXO_ERR_ARG, "exo_policy_set_weights: built for state %d / hidden %d / action %d, got %d / %d / %d", SD, HD, AD,
                             state_dim, hidden_dim, action_dim);
    const float* ptrs[] = {w->zs1_w, w->zs1_b, w->zs2_w, w->zs2_b, w->zs3_w, w->zs3_b, w->l0_w, w->l0_b,
                           w->l1_w,  w->l1_b,  w->l2_w,  w->l2_b,  w->l3_w,  w->l3_b};
    for (const float* q : ptrs)
        if (!q) return exo_set_error(EXO_ERR_ARG, "exo_policy_set_weights: null matrix");
    std::vector<unsigned char> pack(PACK_BYTES);
    unsigned char* d = pack.data();
    // consumption order of the kernel: G1 l0, G2 zs1, G3 zs2, G4 zs3, G5a l1[:,300:600], G5b l1[:,0:300], G6 l2, G7 l3
    // (the l1 bias rides with the a0 half: G5b's operand is the reloaded a0, whose column 300 is the constant 1)
    d += pack_matrix(d, w->l0_w, w->l0_b, HD, SD, 0, SD, N_LO, N_HI, ATOMS_S);
    d += pack_matrix(d, w->zs1_w, w->zs1_b, HD, SD, 0, SD, N_LO, N_HI, ATOMS_S);
    d += pack_matrix(d, w->zs2_w, w->zs2_b, HD, HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix(d, w->zs3_w, w->zs3_b, HD, HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix(d, w->l1_w, nullptr, HD, 2 * HD, HD, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix(d, w->l1_w, w->l1_b, HD, 2 * HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix(d, w->l2_w, w->l2_b, HD, HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix(d, w->l3_w, w->l3_b, AD, HD, 0, HD, OUTP, 0, ATOMS_H);
    if (d != pack.data() + PACK_BYTES) return exo_set_error(EXO_ERR_STATE, "exo_policy_set_weights: pack size mismatch");
    PCK(cudaSetDevice(p->device));
    PCK(cudaMemcpy(p->pack, pack.data(), PACK_BYTES, cudaMemcpyHostToDevice));
    d = pack.data();                 // and the CTA-pair layout of the same matrices
    d += pack_matrix_pair(d, w->l0_w, w->l0_b, HD, SD, 0, SD, N_LO, N_HI, ATOMS_S);
    d += pack_matrix_pair(d, w->zs1_w, w->zs1_b, HD, SD, 0, SD, N_LO, N_HI, ATOMS_S);
    d += pack_matrix_pair(d, w->zs2_w, w->zs2_b, HD, HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix_pair(d, w->zs3_w, w->zs3_b, HD, HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix_pair(d, w->l1_w, nullptr, HD, 2 * HD, HD, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix_pair(d, w->l1_w, w->l1_b, HD, 2 * HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix_pair(d, w->l2_w, w->l2_b, HD, HD, 0, HD, N_LO, N_HI, ATOMS_H);
    d += pack_matrix_pair(d, w->l3_w, w->l3_b, AD, HD, 0, HD, OUTP, 0, ATOMS_H);
    if (d != pack.data() + PACK_BYTES) return exo_set_error(EXO_ERR_STATE, "exo_policy_set_weights: pair pack size mismatch");
    PCK(cudaMemcpy(p->pack2, pack.data(), PACK_BYTES, cudaMemcpyHostToDevice));
    if (int rc = exo_policy_f32_set_weights(p->f32, w)) return rc;
    p->loaded = true;
    return EXO_OK;
}

extern "C" int exo_policy_forward(ExoPolicy* p, const float* obs_dev, int64_t n, float max_action, float* actions_dev, float* zs_dev,
                                  void* stream) {
    if (!p || !obs_dev || !actions_dev || n < 0) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward: bad arguments");
    if (!p->loaded) return exo_set_error(EXO_ERR_STATE, "exo_policy_forward: no weights (exo_policy_set_weights first)");
    if (((uintptr_t)obs_dev & 15) != 0) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward: obs must be 16-byte aligned");
    if (n == 0) return EXO_OK;
    PolicyArgs a;
    a.obs = obs_dev; a.pack = p->pack; a.scratch = p->scratch; a.actions = actions_dev; a.zs = zs_dev;
    a.n = n; a.max_action = max_action; a.dbg = p->dbg;
    const long long tiles = (n + TILE_M - 1) / TILE_M;
    if (tiles > 0x7FFFFFFF) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward: too many observations");
    a.n_tiles = (int)tiles;
    if (p->pair) {
        // clusters of 2 CTAs: an even grid; a CTA whose tile lies past the end works on masked rows to keep its peer company
        const long long want = (tiles + 1) & ~1ll, cap = p->n_sm & ~1;
        a.pack = p->pack2;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(want < cap ? want : cap));
        cfg.blockDim = dim3(N_THREADS);
        cfg.dynamicSmemBytes = SMEM_BYTES;
        cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute attr;
        attr.id = cudaLaunchAttributeClusterDimension;
        attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
        cfg.attrs = &attr; cfg.numAttrs = 1;
        PCK(cudaLaunchKernelEx(&cfg, exo_policy_kernel<true>, a));
    } else {
        const int grid = (int)(tiles < p->n_sm ? tiles : p->n_sm);
        exo_policy_kernel<false><<<grid, N_THREADS, SMEM_BYTES, (cudaStream_t)stream>>>(a);
    }
    PCK(cudaGetLastError());
    p->launches += 1;
    return EXO_OK;
}

extern "C" int exo_policy_forward_fp32(ExoPolicy* p, const float* obs_dev, int64_t n, float max_action, float* actions_dev, float* zs_dev,
                                       void* stream) {
    if (!p || !obs_dev || !actions_dev || n < 0) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward_fp32: bad arguments");
    if (!p->loaded) return exo_set_error(EXO_ERR_STATE, "exo_policy_forward_fp32: no weights (exo_policy_set_weights first)");
    if (((uintptr_t)obs_dev & 15) != 0) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward_fp32: obs must be 16-byte aligned");
    if (n == 0) return EXO_OK;
    if (int rc = exo_policy_f32_forward(p->f32, obs_dev, n, max_action, actions_dev, zs_dev, stream)) return rc;
    p->launches += 1;
    return EXO_OK;
}

// No allocation on this path: argument checks, one kernel launch (exo_policy_fp32.cu, scratch in a static __device__ array).
extern "C" int exo_policy_forward_fp32_noalloc(ExoPolicy* p, const float* obs_dev, int64_t n, float max_action, float* actions_dev,
                                               float* zs_dev, void* stream) {
    if (!p || !obs_dev || !actions_dev || n < 0) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward_fp32_noalloc: bad arguments");
    if (!p->loaded) return exo_set_error(EXO_ERR_STATE, "exo_policy_forward_fp32_noalloc: no weights (exo_policy_set_weights first)");
    if (((uintptr_t)obs_dev & 15) != 0) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward_fp32_noalloc: obs must be 16-byte aligned");
    if (n == 0) return EXO_OK;
    if (int rc = exo_policy_f32_forward_noalloc(p->f32, obs_dev, n, max_action, actions_dev, zs_dev, stream)) return rc;
    p->launches += 1;
    return EXO_OK;
}

extern "C" int exo_policy_profile(ExoPolicy* p, const float* obs_dev, int64_t n, float* actions_dev, int64_t* clocks_host, void* stream) {
    if (!p || !clocks_host) return exo_set_error(EXO_ERR_ARG, "exo_policy_profile: bad arguments");
    long long* d = nullptr;
    PCK(cudaSetDevice(p->device));
    PCK(cudaMalloc((void**)&d, 32 * sizeof(long long)));
    PCK(cudaMemsetAsync(d, 0, 32 * sizeof(long long), (cudaStream_t)stream));
    p->dbg = d;
    const int rc = exo_policy_forward(p, obs_dev, n, 1.0f, actions_dev, nullptr, stream);
    p->dbg = nullptr;
    if (rc) { cudaFree(d); return rc; }
    long long h[32];
    PCK(cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PCK(cudaStreamSynchronize((cudaStream_t)stream));
    PCK(cudaFree(d));
    for (int i = 0; i < 32; ++i) clocks_host[i] = h[i];
    return EXO_OK;
}

extern "C" int64_t exo_policy_launch_count(const ExoPolicy* p) { return p ? p->launches : 0; }
