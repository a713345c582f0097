// exo_reset.cuh — episode (re)initialisation: everything ExoskeletonEnv_train.reset() does
// (Exoskeleton_env.py:473-478 -> initialize_movement :193-254), ONE WARP PER ENVIRONMENT, one kernel.
//
//   exo_seed_kernel    np.random.seed(s) per env: MT19937 init_genrand (Utilities/seed_setting_.py:9)
//   exo_reset_kernel   for each env of the work list (finished envs compacted by the step kernel, or a mask):
//     1. the env's MT19937 key is streamed into shared memory and twisted warp-parallel; the 416 + 9 L 32-bit
//        outputs the episode consumes are tempered into a shared-memory stream, so every later draw is a
//        random-access read and the L-long loops run one sample per lane;
//     2. numpy legacy-RNG draws in the reference's order (a17): tremor magnitude, f1, f2, noise[L], per axis a1, a2,
//        sign[L]; tremor table synthesis (Utilities/generate_parkinson_tremor.py:5-73); symmetric matrix noise
//        (Utilities/domain_randomization_anatomical_matrices.py:4-27); actuator end shifts
//        (Exoskeleton_sim_pybullet.py:98-107); force-range scale (:216-217); counts = 2 (:220); position vectors
//        (:236-250) and the first observation (:252);
//     3. the propagator G = [I 0] * int_0^dt e^{As} ds * [0; I^-1] (fp64 Taylor + scaling and squaring in shared
//        memory), so that each later step replaces scipy's RK45 (calculate_joint_angles.py:5-22, zero initial
//        state every step, env:141,409) by q(dt) = G * T.
//
// All random draws are fp64 and bit-identical to numpy's RandomState: random_sample = (a>>5, b>>6)/2^53 from two
// 32-bit MT outputs, uniform = lo + (hi-lo)*u, choice([-1,1]) = one 32-bit output & 1. Arithmetic that feeds stored
// parameters uses explicit round-to-nearest intrinsics (no FMA contraction) so the values match numpy's
// op-by-op evaluation; sin() and pow() are the only non-IEEE steps (<= 2 ulp from glibc/numpy).
#pragma once
#include "exo_step.cuh"

namespace exo {

#define MT_N 624
#define MT_M 397
#define RST_WARPS 4
#define PD 14

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return y;
}

// Warp-parallel in-place twist of a 624-word key in shared memory. new[i] depends on old[i], old[i+1] and
// (i < 227) old[i+397] / (i >= 227) new[i-227]. Rounds of 128 consecutive words (4 per lane) are independent as long
// as a round's reads complete before its writes: what a round reads of new[] (indices >= 99 below its start) was
// written by earlier rounds, and old[i+1] / old[i+397] lie above everything written so far. Word 623 uses new[0].
// Five rounds with four-way instruction-level parallelism instead of twenty serial 32-word batches. One copy in
// the kernel (__noinline__): the r01d profile showed 33 % of the stall samples waiting on instruction fetch.
static __device__ __noinline__ void mt_twist_warp(uint32_t* s, int lane) {
    for (int b = 0; b < MT_N; b += 128) {
        uint32_t v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = b + 32 * u + lane;
            v[u] = 0;
            if (i < MT_N) {
                const uint32_t cur = s[i], nxt = s[i + 1 < MT_N ? i + 1 : 0];
                const uint32_t y = (cur & 0x80000000u) | (nxt & 0x7fffffffu);
                const int m = i + MT_M < MT_N ? i + MT_M : i + MT_M - MT_N;
                v[u] = s[m] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = b + 32 * u + lane;
            if (i < MT_N) s[i] = v[u];
        }
        __syncwarp();
    }
}

// 32 envs per warp: each lane runs init_genrand for its env, 32 words at a time, through a shared tile so the
// env-major key array is written with coalesced 128 B rows.
static __global__ void __launch_bounds__(32) exo_seed_kernel(uint32_t* mt, int n, const uint32_t* __restrict__ seeds) {
    __shared__ uint32_t tile[32][33];
    const int lane = threadIdx.x, base = blockIdx.x * 32, e = base + lane;
    uint32_t v = e < n ? seeds[e] : 0u;
    for (int chunk = 0; chunk < (MT_N + 31) / 32; chunk++) {
        for (int w = 0; w < 32; w++) {
            const int i = chunk * 32 + w;
            if (i > 0) v = 1812433253u * (v ^ (v >> 30)) + (uint32_t)i;
            tile[lane][w] = v;
        }
        __syncwarp();
        const int i = chunk * 32 + lane;
        for (int r = 0; r < 32 && base + r < n; r++)
            if (i < MT_N) mt[(size_t)(base + r) * EXO_MT_PITCH + i] = tile[r][lane];
        __syncwarp();
    }
    if (e < n) mt[(size_t)e * EXO_MT_PITCH + MT_N] = MT_N;            // position: the first draw twists
}

// seven_by_seven (Utilities/differential_eq_matrices.py:41-69)
static __constant__ double c_I7[49] = {
    0.269, 0, 0, 0.076, 0, 0, -0.014,   0, 0.196, 0.083, 0, -0.002, 0.009, 0,   0, 0.083, 0.079, 0, 0, 0.011, 0,
    0.076, 0, 0, 0.076, 0, 0, -0.012,   0, -0.002, 0, 0, 0.002, 0, 0,           0, 0.009, 0.011, 0, 0, 0.003, 0,
    -0.014, 0, 0, -0.012, 0, 0, 0.003};
static __constant__ double c_D7[49] = {
    0.756, 0.184, 0.020, 0.187, 0, 0, 0,   0.184, 0.383, 0.267, 0, 0, 0, 0,   0.020, 0.267, 0.524, 0, 0, 0, 0,
    0.187, 0, 0, 0.607, 0, 0, 0,           0, 0, 0, 0, 0.021, 0.001, 0.008,   0, 0, 0, 0, 0.001, 0.028, -0.003,
    0, 0, 0, 0, 0.008, -0.003, 0.082};
static __constant__ double c_S7[49] = {
    10.80, 2.626, 0.279, 2.670, 0, 0, 0,   2.626, 5.468, 3.821, 0, 0, 0, 0,   0.279, 3.821, 7.486, 0, 0, 0, 0,
    2.670, 0, 0, 8.670, 0, 0, 0,           0, 0, 0, 0, 0.756, 0.018, 0.291,   0, 0, 0, 0, 0.018, 0.992, -0.099,
    0, 0, 0, 0, 0.291, -0.099, 2.920};
static __constant__ double c_JMAX[7] = {2.5, 5, 10, 5, 5, 0.5, 0.5};    // generate_parkinson_tremor.py:59
// 1/k for the Taylor series of the propagator (a division would take the slow path on every exact zero of the
// padded 16x16 operands: 38 % of the kernel's instructions in the r01c profile)
static __constant__ double c_RCP[16] = {0.0, 1.0, 1.0 / 2, 1.0 / 3, 1.0 / 4, 1.0 / 5, 1.0 / 6, 1.0 / 7, 1.0 / 8, 1.0 / 9, 1.0 / 10,
                                        1.0 / 11, 1.0 / 12, 1.0 / 13, 1.0 / 14, 1.0 / 15};

// legacy random_sample from two consecutive 32-bit outputs of the stream
__device__ __forceinline__ double rs_sample(const uint32_t* strm, int idx) {
    const uint32_t a = strm[idx] >> 5, b = strm[idx + 1] >> 6;
    return __dmul_rn(__dadd_rn(__dmul_rn((double)a, 67108864.0), (double)b), 1.1102230246251565e-16);   // / 2^53, exact
}
__device__ __forceinline__ double rs_uniform(const uint32_t* strm, int idx, double lo, double hi) {   // lo + (hi-lo)*u
    return __dadd_rn(lo, __dmul_rn(__dsub_rn(hi, lo), rs_sample(strm, idx)));
}

__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------------------------------------------------
// MT19937 stream consumption, warp-cooperative. `pos` is the next unread word of the shared key.
// ---------------------------------------------------------------------------------------------------------
// next n tempered outputs -> dst[0..n)
static __device__ __noinline__ void mt_fill_words(uint32_t* key, int& pos, uint32_t* dst, int n, int lane) {
    for (int done = 0; done < n;) {
        if (pos >= MT_N) { mt_twist_warp(key, lane); pos = 0; }
        const int take = min(MT_N - pos, n - done);
        for (int i = lane; i < take; i += 32) dst[done + i] = mt_temper(key[pos + i]);
        done += take; pos += take;
    }
}
// next n outputs -> bit k of bits[] = (output k) & 1   (np.random.choice([-1, 1]) draws, tremor:70)
static __device__ __noinline__ void mt_fill_bits(uint32_t* key, int& pos, uint32_t* bits, int n, int lane) {
    for (int done = 0; done < n;) {
        if (pos >= MT_N) { mt_twist_warp(key, lane); pos = 0; }
        const int take = min(min(MT_N - pos, n - done), 32 - (done & 31));      // never straddles a 32-bit word
        const uint32_t y = lane < take ? mt_temper(key[pos + lane]) : 0u;
        const uint32_t b = __ballot_sync(0xffffffffu, (y & 1u) != 0u) << (done & 31);
        if (lane == 0) { if ((done & 31) == 0) bits[done >> 5] = b; else bits[done >> 5] |= b; }
        done += take; pos += take;
    }
}
// consume n outputs without using them (the state still has to advance)
static __device__ __noinline__ void mt_skip(uint32_t* key, int& pos, int n, int lane) {
    while (n > 0) {
        if (pos >= MT_N) { mt_twist_warp(key, lane); pos = 0; }
        const int take = min(MT_N - pos, n);
        pos += take; n -= take;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Propagator (one warp, fp64 on the tensor cores: mma.sync m8n8k4 f64 = DMMA).
//   A = [[0, 1], [-I^-1 S, -I^-1 D]] (14x14),  B = [0; I^-1] (14x7)
//   E = e^{A h}, Phi = sum_k (A h)^k/(k+1)!  (Taylor, degree 14, h = dt/2^s with ||A h||_inf <= 1/4)
//   F = h * Phi * B, then s times: F <- (E + 1) F ; E <- E E.  G = rows 0..6 of F.
// Matrices are padded to 16x16 (pitch PP doubles, pads stay exactly zero); E, Phi, P and F live in registers as
// accumulator fragments, shared memory only converts accumulator layout into operand layout between products.
// Fragment maps (PTX ISA, m8n8k4 .f64): A(8x4) row = lane/4, col = lane%4 | B(4x8) row = lane%4, col = lane/4 |
// C(8x8) row = lane/4, cols = 2*(lane%4) + {0,1}.
// ids: I, D, S (3 x 49, shared). sm: 3 * 16 * PP + 98 doubles of scratch.
// ---------------------------------------------------------------------------------------------------------
#define PP 20      // pitch (doubles): bank = (8 row + 2 col) mod 32 -> conflict-free operand loads per half-warp
#define PROP_SCRATCH_DOUBLES (3 * 16 * PP + 98)

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
// accumulator fragments of a 16x16 matrix: c[(rt*2+ct)*2 + i] = M[rt*8 + lane/4][ct*8 + 2*(lane%4) + i]
__device__ __forceinline__ void frag_store_c(double* M, const double* c, int lane) {
    const int r = lane >> 2, q = (lane & 3) * 2;
#pragma unroll
    for (int rt = 0; rt < 2; rt++)
#pragma unroll
        for (int ct = 0; ct < 2; ct++)
            *reinterpret_cast<double2*>(&M[(rt * 8 + r) * PP + ct * 8 + q]) = make_double2(c[(rt * 2 + ct) * 2], c[(rt * 2 + ct) * 2 + 1]);
}
__device__ __forceinline__ void frag_load_c(const double* M, double* c, int lane) {
    const int r = lane >> 2, q = (lane & 3) * 2;
#pragma unroll
    for (int rt = 0; rt < 2; rt++)
#pragma unroll
        for (int ct = 0; ct < 2; ct++) {
            const double2 v = *reinterpret_cast<const double2*>(&M[(rt * 8 + r) * PP + ct * 8 + q]);
            c[(rt * 2 + ct) * 2] = v.x; c[(rt * 2 + ct) * 2 + 1] = v.y;
        }
}
// left-operand fragments: a[rt*4 + kt] = M[rt*8 + lane/4][kt*4 + lane%4]
__device__ __forceinline__ void frag_load_a(const double* M, double* a, int lane) {
#pragma unroll
    for (int rt = 0; rt < 2; rt++)
#pragma unroll
        for (int kt = 0; kt < 4; kt++) a[rt * 4 + kt] = M[(rt * 8 + (lane >> 2)) * PP + kt * 4 + (lane & 3)];
}
// right-operand fragments: b[kt*NCT + ct] = M[kt*4 + lane%4][ct*8 + lane/4]
template <int NCT>
__device__ __forceinline__ void frag_load_b(const double* M, double* b, int lane) {
#pragma unroll
    for (int kt = 0; kt < 4; kt++)
#pragma unroll
        for (int ct = 0; ct < NCT; ct++) b[kt * NCT + ct] = M[(kt * 4 + (lane & 3)) * PP + ct * 8 + (lane >> 2)];
}
// c (16 x 8*NCT, accumulator fragments) += a (16x16) * b (16 x 8*NCT)
template <int NCT>
__device__ __forceinline__ void frag_mma(double* c, const double* a, const double* b) {
#pragma unroll
    for (int rt = 0; rt < 2; rt++)
#pragma unroll
        for (int ct = 0; ct < NCT; ct++)
#pragma unroll
            for (int kt = 0; kt < 4; kt++) dmma(c[(rt * NCT + ct) * 2], c[(rt * NCT + ct) * 2 + 1], a[rt * 4 + kt], b[kt * NCT + ct]);
}

// Returns this lane's accumulator fragments of F (16x8): f[rt*2 + i] = F[rt*8 + lane/4][2*(lane%4) + i].
static __device__ __noinline__ void warp_propagator(const double* ids, double* sm, double dt, int lane, double* f) {
    double* M0 = sm;                 // A h, later B = [0; I^-1]
    double* M1 = M0 + 16 * PP;       // P / Phi / E in operand layout
    double* M2 = M1 + 16 * PP;       // F in operand layout
    double* LU = M2 + 16 * PP;       // 49: I, factorised in place (Gauss-Jordan -> I^-1)
    double* Iinv = LU + 49;          // 49
    // ---- I^-1 by Gauss-Jordan with partial pivoting (lane-parallel over columns of [I | 1])
    for (int i = lane; i < 49; i += 32) { LU[i] = ids[i]; Iinv[i] = (i / 7 == i % 7) ? 1.0 : 0.0; }
    __syncwarp();
    for (int k = 0; k < 7; k++) {
        int p = k; double best = fabs(LU[k * 7 + k]);
        for (int i = k + 1; i < 7; i++) { const double v = fabs(LU[i * 7 + k]); if (v > best) { best = v; p = i; } }
        __syncwarp();
        if (p != k && lane < 14) {
            double* M = lane < 7 ? LU : Iinv; const int j = lane % 7;
            const double t = M[k * 7 + j]; M[k * 7 + j] = M[p * 7 + j]; M[p * 7 + j] = t;
        }
        __syncwarp();
        const double d = LU[k * 7 + k];
        __syncwarp();
        if (lane < 14) { double* M = lane < 7 ? LU : Iinv; M[k * 7 + lane % 7] /= d; }
        __syncwarp();
        double fk[7];
#pragma unroll
        for (int i = 0; i < 7; i++) fk[i] = LU[i * 7 + k];
        __syncwarp();
        if (lane < 14) {
            double* M = lane < 7 ? LU : Iinv; const int j = lane % 7;
            const double pk = M[k * 7 + j];
#pragma unroll
            for (int i = 0; i < 7; i++) if (i != k) M[i * 7 + j] -= fk[i] * pk;
        }
        __syncwarp();
    }
    // ---- A (unscaled, zero-padded to 16 x PP) into M0, row-sum norm
    for (int idx = lane; idx < 16 * PP; idx += 32) {
        const int i = idx / PP, j = idx - i * PP;
        double v = 0;
        if (i < 7) v = (j == i + 7) ? 1.0 : 0.0;
        else if (i < PD && j < PD) {
            const double* M = ids + (j < 7 ? 98 : 49);   // S for the q block, D for the qdot block
            const int jj = j < 7 ? j : j - 7;
            double acc = 0;
#pragma unroll
            for (int k = 0; k < 7; k++) acc = fma(Iinv[(i - 7) * 7 + k], M[k * 7 + jj], acc);
            v = -acc;
        }
        M0[idx] = v;
    }
    __syncwarp();
    double nrm = 0;
    if (lane < PD) { for (int j = 0; j < PD; j++) nrm += fabs(M0[lane * PP + j]); }
    for (int o = 16; o > 0; o >>= 1) nrm = fmax(nrm, __shfl_xor_sync(0xffffffffu, nrm, o));
    nrm *= dt;
    int s = 0;
    while (nrm > 0.25 && s < 40) { nrm *= 0.5; s++; }
    const double h = ldexp(dt, -s);
    for (int idx = lane; idx < 16 * PP; idx += 32) M0[idx] *= h;
    __syncwarp();
    // ---- Taylor: P_1 = A h; P_k = P_{k-1} (A h) / k; E = 1 + sum P_k; Phi = 1 + sum P_k / (k+1)
    double bA[8], P[8], E[8], Phi[8];
    frag_load_b<2>(M0, bA, lane);
    frag_load_c(M0, P, lane);
#pragma unroll
    for (int t = 0; t < 8; t++) {
        const int row = (t >> 2) * 8 + (lane >> 2), col = ((t >> 1) & 1) * 8 + (lane & 3) * 2 + (t & 1);
        const double id = (row == col) ? 1.0 : 0.0;
        E[t] = id + P[t]; Phi[t] = id + P[t] * 0.5;
    }
    for (int k = 2; k <= 14; k++) {
        frag_store_c(M1, P, lane);
        __syncwarp();
        double a[8];
        frag_load_a(M1, a, lane);
#pragma unroll
        for (int t = 0; t < 8; t++) P[t] = 0.0;
        frag_mma<2>(P, a, bA);
        const double inv = c_RCP[k], inv1 = c_RCP[k + 1];
#pragma unroll
        for (int t = 0; t < 8; t++) { P[t] *= inv; E[t] += P[t]; Phi[t] = fma(P[t], inv1, Phi[t]); }
        __syncwarp();
    }
    // ---- F = h * Phi * B,  B[7 + r][c] = I^-1[r][c]  (16 x 8 operand in M0)
    frag_store_c(M1, Phi, lane);
    for (int idx = lane; idx < 16 * PP; idx += 32) {
        const int i = idx / PP, j = idx - i * PP;
        M0[idx] = (i >= 7 && i < PD && j < 7) ? Iinv[(i - 7) * 7 + j] : 0.0;
    }
    __syncwarp();
    {
        double a[8], b[4];
        frag_load_a(M1, a, lane);
        frag_load_b<1>(M0, b, lane);
#pragma unroll
        for (int t = 0; t < 4; t++) f[t] = 0.0;
        frag_mma<1>(f, a, b);
#pragma unroll
        for (int t = 0; t < 4; t++) f[t] *= h;
    }
    __syncwarp();
    // ---- undo the scaling: F <- F + E F ; E <- E E
    for (int it = 0; it < s; it++) {
        frag_store_c(M1, E, lane);
        {   // F (16 x 8) in operand layout
            const int r = lane >> 2, q = (lane & 3) * 2;
            *reinterpret_cast<double2*>(&M2[r * PP + q]) = make_double2(f[0], f[1]);
            *reinterpret_cast<double2*>(&M2[(8 + r) * PP + q]) = make_double2(f[2], f[3]);
        }
        __syncwarp();
        double a[8], b[4];
        frag_load_a(M1, a, lane);
        frag_load_b<1>(M2, b, lane);
        frag_mma<1>(f, a, b);
        if (it + 1 < s) {
            double bE[8];
            frag_load_b<2>(M1, bE, lane);
#pragma unroll
            for (int t = 0; t < 8; t++) E[t] = 0.0;
            frag_mma<2>(E, a, bE);
        }
        __syncwarp();
    }
}

// Shared memory per warp (bytes): region A = stream words 4*(416 + 2*TL) | sign bits 7*4*ceil(TL/32) |
// waves 2*8*TL (sin w1 t, sin w2 t; first the MT key, 2496 B, dead before the waves are written); the whole of
// region A is aliased by the propagator scratch (which runs after all of it is dead) | ids 148*8 | small 160*8.
// 12.5 KB per warp at table_len 347 -> 4 blocks of 4 warps per SM.
__host__ __device__ inline size_t reset_stream_words(int table_len) { return (size_t)((416 + 2 * table_len + 3) & ~3); }
__host__ __device__ inline size_t reset_sign_words(int table_len) { return (size_t)((((table_len + 31) / 32) * 7 + 3) & ~3); }
__host__ __device__ inline size_t reset_region_a(int table_len) {
    size_t w = 2 * 8 * (size_t)((table_len + 1) & ~1);
    if (w < 2496) w = 2496;
    size_t a = 4 * reset_stream_words(table_len) + 4 * reset_sign_words(table_len) + w;
    const size_t scratch = PROP_SCRATCH_DOUBLES * 8;
    return ((a > scratch ? a : scratch) + 15) & ~(size_t)15;
}
__host__ __device__ inline size_t reset_smem_per_warp(int table_len) {
    return reset_region_a(table_len) + 148 * 8 + 160 * 8;                 // multiple of 16
}


// ---------------------------------------------------------------------------------------------------------
// exo_gen_kernel: episode parameters of the NEXT episode of each env of the work list, into the env's spare slot
// (the slot bit of `counts` names the slot in use). Runs one episode ahead of its consumer, on a side stream.
// ---------------------------------------------------------------------------------------------------------
template <typename R, int NAX>
__global__ void __launch_bounds__(RST_WARPS * 32)
exo_gen_kernel(EnvArrays<R> S, ResetConst K, const uint8_t* __restrict__ mask) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int NAXV = (NAX + 3) / 4;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int TL = (S.table_len + 1) & ~1, st = S.stride;   // wave pitch (even: keeps every region 16-byte aligned)
    unsigned char* wsm = smem_raw + (size_t)wid * reset_smem_per_warp(S.table_len);
    double* scratch = reinterpret_cast<double*>(wsm);                      // propagator scratch aliases region A
    uint32_t* strm = reinterpret_cast<uint32_t*>(wsm);                     // compact stream: head | 7 x 4 amplitude words | tail
    uint32_t* sbits = strm + reset_stream_words(S.table_len);              // [7][SW] sign bits
    const int SW = (S.table_len + 31) / 32;
    double* wav = reinterpret_cast<double*>(sbits + reset_sign_words(S.table_len));   // [2][TL]: sin w1 t, sin w2 t
    uint32_t* key = reinterpret_cast<uint32_t*>(wav);                      // [624] MT key: dead before wav is written
    double* ids = reinterpret_cast<double*>(wsm + reset_region_a(S.table_len));   // [148]
    double* sml = ids + 148;                                               // [160] small per-env scratch
    // sml layout: 0..13 a1[7],a2[7] | 16..57 shift[42] | 58,59 fmax

    const int nwarps = blockDim.x >> 5;                                    // 1 .. RST_WARPS warps per block (launcher's choice)
    const int total_warps = gridDim.x * nwarps;
    const int gw = blockIdx.x * nwarps + wid;
    const int* list = nullptr;
    int work = S.n;
    if (K.list >= 0) { work = S.fin_count[K.list]; list = S.fin_list + (size_t)K.list * st; }

    for (int wi = gw; wi < work; wi += total_warps) {
        const int e = list ? list[wi] : wi;
        if (!list && mask != nullptr && mask[e] == 0) continue;
        const int meta = S.meta[e];
        const int L = meta >> 16;
        const int slot = ((S.counts[e] >> EXO_SLOT_SHIFT) & 1) ^ 1;        // the spare slot
        V4<R>* tt = S.tt + (size_t)slot * S.tt_slot;
        V4<R>* ep = S.ep + (size_t)slot * S.ep_slot;
        V4<R>* Gs = S.G + (size_t)slot * S.G_slot;
        double* IDS = S.IDS + ((size_t)slot * S.n + e) * EXO_IDS_PITCH;
        // ---- 1. MT19937: key -> shared, generate the episode's outputs into the stream
        uint32_t* gkey = S.mt + (size_t)e * EXO_MT_PITCH;
        for (int i = lane; i < MT_N; i += 32) key[i] = gkey[i];
        int pos = (int)gkey[MT_N];
        __syncwarp();
        // draw order of initialize_movement: [ctor rand()] mag f1 f2 noise[L] | per axis a1 a2 sign[L] | I D S
        // noise (3 x 49) shifts (42) force scales (2). Full words are kept for the fp64 draws, one bit per sign draw.
        const int o_mag = 0, o_f1 = 2, o_f2 = 4, o_noise = 6;
        const int o_axis = o_noise + 2 * L;                     // 7 x (a1, a2) = 4 words each
        const int o_mat = o_axis + 28;                          // I, D, S: 49 x 2 each
        const int o_shift = o_mat + 294, o_fsh = o_shift + 84, o_fel = o_fsh + 2;
        mt_skip(key, pos, K.ctor_skip, lane);
        mt_fill_words(key, pos, strm, o_axis, lane);
        for (int j = 0; j < 7; j++) {
            mt_fill_words(key, pos, strm + o_axis + 4 * j, 4, lane);
            if (j < NAX && ((meta >> j) & 1)) mt_fill_bits(key, pos, sbits + j * SW, L, lane);
            else mt_skip(key, pos, L, lane);
        }
        mt_fill_words(key, pos, strm + o_mat, 382, lane);
        __syncwarp();
        for (int i = lane; i < MT_N; i += 32) gkey[i] = key[i];
        if (lane == 0) gkey[MT_N] = (uint32_t)pos;
        __syncwarp();                                           // the key's shared copy is dead: its space becomes wav
        // ---- tremor magnitude (:198-199)
        const double mag = __dadd_rn(K.amp_lo, __dmul_rn(rs_sample(strm, o_mag), __dsub_rn(K.amp_hi, K.amp_lo)));
        // ---- generate_tremor_frequency (tremor:5-21): f1, f2, white noise, time base, the two harmonics
        const double f1 = rs_uniform(strm, o_f1, K.f1_lo, K.f1_hi);
        const double f2 = rs_uniform(strm, o_f2, K.f2_lo, K.f2_hi);
        const double stop = __dmul_rn((double)L, K.dt), tstep = __ddiv_rn(stop, (double)(L - 1));   // np.linspace(0, L*dt, L)
        const double w1c = __dmul_rn(__dmul_rn(2.0, 3.141592653589793), f1), w2c = __dmul_rn(__dmul_rn(2.0, 3.141592653589793), f2);
        for (int kk = lane; kk < 2 * L; kk += 32) {            // one sin() call site: both harmonics in one loop
            const int hsel = kk >= L, k = hsel ? kk - L : kk;
            const double t = (k == L - 1) ? stop : __dmul_rn((double)k, tstep);
            wav[hsel * TL + k] = sin(__dmul_rn(hsel ? w2c : w1c, t));
        }
        // ---- per-axis amplitudes (tremor:24-28): lanes 0..13 take one pow each
        if (lane < 14) {
            const int j = lane >> 1, o = o_axis + 4 * j + 2 * (lane & 1);
            const double db = (lane & 1) ? rs_uniform(strm, o, -20.0, -10.0) : rs_uniform(strm, o, -5.0, 0.0);
            sml[(lane & 1) * 7 + j] = pow(10.0, __ddiv_rn(db, 20.0));
        }
        // ---- add_symmetric_noise on I, D, S (:208-210): (n + n^T)/2, M + n*M
        for (int idx = lane; idx < 147; idx += 32) {
            const int m = idx / 49, ij = idx - m * 49, i = ij / 7, j = ij - i * 7;
            const double* M0 = (m == 0) ? c_I7 : (m == 1 ? c_D7 : c_S7);
            const double nij = rs_uniform(strm, o_mat + m * 98 + 2 * (i * 7 + j), -K.noise_frac, K.noise_frac);
            const double nji = rs_uniform(strm, o_mat + m * 98 + 2 * (j * 7 + i), -K.noise_frac, K.noise_frac);
            const double sij = __ddiv_rn(__dadd_rn(nij, nji), 2.0);
            const double v = __dadd_rn(M0[ij], __dmul_rn(sij, M0[ij]));
            ids[idx] = v;
            IDS[idx] = v;
        }
        // ---- actuator end shifts (sim:98-107) and force range scale (:216-217)
        for (int i = lane; i < 42; i += 32) sml[16 + i] = rs_uniform(strm, o_shift + 2 * i, -K.shift_range, K.shift_range);
        if (lane < 2) {
            const double lo = __dsub_rn(1.0, K.act_range), hi = __dadd_rn(1.0, K.act_range);
            sml[58 + lane] = __dmul_rn(lane == 0 ? K.fmax_sh0 : K.fmax_el0, rs_uniform(strm, lane == 0 ? o_fsh : o_fel, lo, hi));
        }
        __syncwarp();
        // ---- min/max of the raw wave per axis (tremor:64), then the normalised, sign-flipped table (tremor:64-70)
        double a1[NAX], a2[NAX], mn[NAX], mx[NAX];
#pragma unroll
        for (int j = 0; j < NAX; j++) { a1[j] = sml[j]; a2[j] = sml[7 + j]; mn[j] = INFINITY; mx[j] = -INFINITY; }
        for (int k = lane; k < L; k += 32) {
            const double w1 = wav[k], w2 = wav[TL + k], nz = __dmul_rn(rs_sample(strm, o_noise + 2 * k), 0.001);   // tremor:12
#pragma unroll
            for (int j = 0; j < NAX; j++) {
                const double gk = __dadd_rn(__dadd_rn(__dmul_rn(a1[j], w1), __dmul_rn(a2[j], w2)), nz);
                mn[j] = fmin(mn[j], gk); mx[j] = fmax(mx[j], gk);
            }
        }
        // (g - min) / (max - min) as a correctly rounded division from ONE reciprocal per axis (Markstein: y = RN(1/d),
        // q = RN(x y), r = x - d q exactly, RN(q + r y) = RN(x / d)) instead of a ~25-instruction division per sample
        double dd[NAX], ry[NAX], jm[NAX];
#pragma unroll
        for (int j = 0; j < NAX; j++) {
            mn[j] = warp_min(mn[j]); mx[j] = warp_max(mx[j]);
            dd[j] = __dsub_rn(mx[j], mn[j]); ry[j] = __drcp_rn(dd[j]); jm[j] = __dmul_rn(c_JMAX[j], mag);
        }
        for (int k = lane; k < L; k += 32) {
            const double w1 = wav[k], w2 = wav[TL + k], nz = __dmul_rn(rs_sample(strm, o_noise + 2 * k), 0.001);
            R row[4 * NAXV];
#pragma unroll
            for (int j = 0; j < 4 * NAXV; j++) {
                double v = 0.0;
                if (j < NAX && ((meta >> j) & 1)) {
                    const double gk = __dadd_rn(__dadd_rn(__dmul_rn(a1[j], w1), __dmul_rn(a2[j], w2)), nz);
                    const double x = __dmul_rn(2.0, __dsub_rn(gk, mn[j]));
                    const double q0 = __dmul_rn(x, ry[j]);
                    const double q = __fma_rn(__fma_rn(-dd[j], q0, x), ry[j], q0);
                    v = __dmul_rn(__dadd_rn(-1.0, q), jm[j]);
                    if (isnan(v) || isinf(v)) v = 0.0;                                                  // tremor:67
                    if (((sbits[j * SW + (k >> 5)] >> (k & 31)) & 1u) == 0) v = -v;                     // choice([-1, 1]) (tremor:70)
                }
                row[j] = (R)v;
            }
#pragma unroll
            for (int h = 0; h < NAXV; h++) st4(tt + ((size_t)k * st + e) * NAXV + h, row + 4 * h);
        }
        // ---- per-episode parameter record: 42 shifts + 2 force maxima, one 16-byte vector per lane
        if (lane < NV_EP) {
            R pv[4];
#pragma unroll
            for (int c4 = 0; c4 < 4; c4++) { const int w = lane * 4 + c4; pv[c4] = (R)(w < 42 ? sml[16 + w] : sml[58 + (w - 42)]); }
            st4(ep + (size_t)lane * st + e, pv);
        }
        if (lane == 0) S.mag[(size_t)slot * st + e] = mag;
        __syncwarp();
        // ---- 3. propagator from this episode's I, D, S on the fp64 tensor cores; scratch aliases region A (dead by now)
        {
            double f[4];
            warp_propagator(ids, scratch, K.dt, lane, f);
            const int row = lane >> 2, col = (lane & 3) * 2;               // rows 0..6 of F are in the first row tile
            if (row < 7) {
                *exo_word(Gs, exo_g_word(row, col), e, st) = (R)f[0];
                if (col + 1 < 7) *exo_word(Gs, exo_g_word(row, col + 1), e, st) = (R)f[1];
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------------
// Beginning of an episode on already generated parameters (Exoskeleton_env.py:220-252): flip to the spare slot,
// counts = 2, the two position reads of the unchanged link-state cache (:230-250) and the first observation (:252).
// One thread per env; shared by exo_begin_kernel (explicit reset()) and the step kernel's auto-reset path, which
// passes the joint vector / elbow reference it already holds and a staging tile for the observation.
//   qs[5]   joint vector behind the link-state cache at this moment;  el_old  act_elbow_reference_positions as last read
//   qpl     EXO_MOTOR_FULL: prismatic positions behind the cache (else nullptr)
//   obA/obB the 80-float observation as 20 float4 chunks: chunks 0..10 -> obA[0..10], chunks 11..19 -> obB[0..8]
//   st_ep / st_tt  fp32 step kernel: the 11 parameter vectors / 3 tremor rows of the new slot, already staged in shared
//           memory by asynchronous copies issued at load time (st_ep may alias obA: it is read before obA is written);
//           nullptr: read them from global memory here
// ---------------------------------------------------------------------------------------------------------
template <typename R, int NAX>
__device__ __forceinline__ void begin_episode(const EnvArrays<R>& S, const ChainConst<R>& C, int e, int new_slot,
                                           const R* qs, const R* el_old, const R* qpl, float4* obA, float4* obB,
                                           const float4* st_ep = nullptr, const float4* st_tt = nullptr) {
    constexpr int NAXV = (NAX + 3) / 4;
    const int st = S.stride;
    const V4<R>* ep = S.ep + (size_t)new_slot * S.ep_slot;
    const V4<R>* tt = S.tt + (size_t)new_slot * S.tt_slot;
    // every global load of this path is issued before the first use (the 11 parameter vectors and the three tremor rows
    // fly while the chain is evaluated), and every store comes after the last load: with loads interleaved between the
    // stores the 21 shift reads were 21 exposed memory latencies (~10 us on the step's critical path, r02g)
    R sh[4 * NV_EP], rows[3][4];
    if (sizeof(R) == 4 && st_ep != nullptr) {
#pragma unroll
        for (int v = 0; v < NV_EP; v++) { const float4 t = st_ep[v]; sh[4 * v] = (R)t.x; sh[4 * v + 1] = (R)t.y; sh[4 * v + 2] = (R)t.z; sh[4 * v + 3] = (R)t.w; }
#pragma unroll
        for (int k = 0; k < 3; k++) { const float4 t = st_tt[k]; rows[k][0] = (R)t.x; rows[k][1] = (R)t.y; rows[k][2] = (R)t.z; rows[k][3] = (R)t.w; }
    } else {
#pragma unroll
        for (int v = 0; v < NV_EP; v++) ld4_ord(ep + (size_t)v * st + e, sh + 4 * v);
#pragma unroll
        for (int k = 0; k < 3; k++) ld4_ord(tt + ((size_t)k * st + e) * NAXV, rows[k]);
    }
    float ob[80];
    ChainPose<R> P;
    chain_fk(C, qs, P);
#pragma unroll
    for (int i = 0; i < 14; i++) ob[i] = 0.0f;                                        // force history zeroed (:195)
#pragma unroll
    for (int k = 0; k < 3; k++)                                                       // tremor rows 0..2 (:502-513)
#pragma unroll
        for (int j = 0; j < 4; j++)               // same arithmetic as the step kernel's observation (fp32: reciprocal multiply)
            ob[14 + k * 4 + j] = sizeof(R) == 8 ? (float)(rows[k][j] / (j == 3 ? R(5) : R(10))) : (float)(rows[k][j] * (j == 3 ? R(0.2) : R(0.1)));
    R pv[21];
#pragma unroll
    for (int i = 0; i < 7; i++) {
        R a2[3], off[3];
#pragma unroll
        for (int k = 0; k < 3; k++) off[k] = i < 2 ? C.off_alk[i][k] : C.off_hum[i][k];
        if (qpl != nullptr) {                  // EXO_MOTOR_FULL: the insertion-end dummy may have drifted along its prismatic axis
            const R qp = qpl[i < 2 ? 7 + i : i];
#pragma unroll
            for (int k = 0; k < 3; k++) off[k] += (i < 2 ? C.pax_alk[i][k] : C.pax_hum[i][k]) * qp;
        }
        if (i < 2) affine(P.R4, P.p4, off, a2);
        else affine(P.R2, P.p2, off, a2);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            a2[k] += sh[W_SHIFT + (2 * i + 1) * 3 + k];
            const R prev = a2[k] - (i < 2 ? el_old[k] : C.sh_ref[k]);      // stale elbow ref on the first read (:237)
            const R cur = a2[k] - (i < 2 ? P.elref[k] : C.sh_ref[k]);      // refreshed before the second (:238,249)
            pv[i * 3 + k] = cur;
            ob[26 + i * 3 + k] = (float)prev; ob[47 + i * 3 + k] = (float)cur;
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        ob[68 + k] = (float)C.sh_ref[k]; ob[71 + k] = (float)P.elref[k];
        ob[74 + k] = (float)C.sh_ref[k]; ob[77 + k] = (float)P.elref[k];
    }
    {   // words 10..33 of the state record = elref[3] | posvec[21]: vectors 3..7 whole, the four words around them singly
        static_assert(W_ELREF == 10 && W_POSVEC == 13, "state layout");
        V4<R>* sv = S.state + e;
        sv[(size_t)2 * st].v[2] = P.elref[0]; sv[(size_t)2 * st].v[3] = P.elref[1];
        const R v3[4] = {P.elref[2], pv[0], pv[1], pv[2]};
        st4(sv + (size_t)3 * st, v3);
#pragma unroll
        for (int v = 4; v < 8; v++) st4(sv + (size_t)v * st, pv + (4 * v - 13));
        sv[(size_t)8 * st].v[0] = pv[19]; sv[(size_t)8 * st].v[1] = pv[20];
    }
    S.ret[e] = R(0);
    S.counts[e] = 2 | (new_slot << EXO_SLOT_SHIFT);                                   // counts = 2 (:220)
#pragma unroll
    for (int ch = 0; ch < 11; ch++) obA[ch] = make_float4(ob[4 * ch], ob[4 * ch + 1], ob[4 * ch + 2], ob[4 * ch + 3]);
#pragma unroll
    for (int ch = 11; ch < 20; ch++) obB[ch - 11] = make_float4(ob[4 * ch], ob[4 * ch + 1], ob[4 * ch + 2], ob[4 * ch + 3]);
}

// explicit env.reset() (mask / all), or the envs of a work list (auto-reset when the flip cannot be fused into the step)
template <typename R, int NAX>
__global__ void __launch_bounds__(64)
exo_begin_kernel(EnvArrays<R> S, ChainConst<R> C, ResetConst K, const uint8_t* __restrict__ mask, float* __restrict__ obs) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x, st = S.stride;
    const int lane = threadIdx.x & 31;
    int e = idx;
    bool act;
    if (K.list >= 0) { act = idx < S.fin_count[K.list]; if (act) e = S.fin_list[(size_t)K.list * st + idx]; }
    else act = idx < S.n && (mask == nullptr || mask[idx] != 0);
    if (act) {
        const int new_slot = ((S.counts[e] >> EXO_SLOT_SHIFT) & 1) ^ 1;
        R qs[5], el_old[3], qpl[EXO_NPRIS];
#pragma unroll
        for (int j = 0; j < 5; j++) qs[j] = *exo_word(S.state, (K.lag ? W_QLAG : W_Q) + j, e, st);
#pragma unroll
        for (int k = 0; k < 3; k++) el_old[k] = *exo_word(S.state, W_ELREF + k, e, st);
        if (K.motor_full) {
#pragma unroll
            for (int d = 0; d < EXO_NPRIS; d++) qpl[d] = *exo_word(S.dyn, (K.lag ? W_QPLAG : W_QP) + d, e, st);
        }
        float4 oa[11], obb[9];
        begin_episode<R, NAX>(S, C, e, new_slot, qs, el_old, K.motor_full ? qpl : nullptr, oa, obb);
        if (obs != nullptr) {
            float4* dst = reinterpret_cast<float4*>(obs + (size_t)e * 80);
#pragma unroll
            for (int ch = 0; ch < 11; ch++) dst[ch] = oa[ch];
#pragma unroll
            for (int ch = 0; ch < 9; ch++) dst[11 + ch] = obb[ch];
        }
    }
    if (K.append >= 0) {                       // the envs started here need their (new) spare slot refilled
        const unsigned m = __ballot_sync(0xffffffffu, act);
        if (m) {
            int base = 0;
            if (lane == 0) base = atomicAdd(S.fin_count + K.append, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (act) S.fin_list[(size_t)K.append * st + base + __popc(m & ((1u << lane) - 1u))] = e;
        }
    }
}

}  // namespace exo
