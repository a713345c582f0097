// exo_reset.cuh — episode (re)initialisation kernels: everything ExoskeletonEnv_train.reset() does
// (Exoskeleton_env.py:473-478 -> initialize_movement :193-254) for the envs selected by a mask.
//
//   exo_seed_kernel        np.random.seed(s) per env: MT19937 init_genrand (Utilities/seed_setting_.py:9)
//   exo_reset_kernel       numpy legacy-RNG draws in the reference's order, tremor table synthesis
//                          (Utilities/generate_parkinson_tremor.py:5-73), symmetric matrix noise
//                          (Utilities/domain_randomization_anatomical_matrices.py:4-27), actuator end shifts
//                          (Exoskeleton_sim_pybullet.py:98-107), force-range scale (:216-217), counts = 2 (:220),
//                          position vectors (:236-250) and the first observation (:252)
//   exo_propagator_kernel  one warp per env: G = [I 0] * int_0^dt e^{As} ds * [0; I^-1] (fp64, scaling and
//                          squaring), so that each later step replaces scipy's RK45 (calculate_joint_angles.py:5-22,
//                          zero initial state every step, env:141,409) by q(dt) = G * T.
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

// In-place twist of one env's 624-word key (global memory, stride-N layout -> coalesced across the warp).
__device__ inline void mt_twist(uint32_t* s, int stride) {
#define MT_AT(i) s[(size_t)(i) * stride]
    uint32_t cur = MT_AT(0);
    uint32_t y, nxt;
    int i = 0;
    uint32_t first_new = 0;
    for (; i < MT_N - MT_M; i++) {
        nxt = MT_AT(i + 1);
        y = (cur & 0x80000000u) | (nxt & 0x7fffffffu);
        const uint32_t v = MT_AT(i + MT_M) ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        MT_AT(i) = v;
        if (i == 0) first_new = v;
        cur = nxt;
    }
    for (; i < MT_N - 1; i++) {
        nxt = MT_AT(i + 1);
        y = (cur & 0x80000000u) | (nxt & 0x7fffffffu);
        MT_AT(i) = MT_AT(i + (MT_M - MT_N)) ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        cur = nxt;
    }
    y = (cur & 0x80000000u) | (first_new & 0x7fffffffu);
    MT_AT(MT_N - 1) = MT_AT(MT_M - 1) ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
#undef MT_AT
}

struct Mt {
    uint32_t* s;
    int stride, pos;
    __device__ __forceinline__ uint32_t next32() {
        if (pos >= MT_N) { mt_twist(s, stride); pos = 0; }
        uint32_t y = s[(size_t)pos * stride];
        pos++;
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    __device__ __forceinline__ double sample() {                  // legacy random_sample / rand()
        const uint32_t a = next32() >> 5, b = next32() >> 6;
        return __ddiv_rn(__dadd_rn(__dmul_rn((double)a, 67108864.0), (double)b), 9007199254740992.0);
    }
    __device__ __forceinline__ double uniform(double lo, double hi) {  // lo + (hi-lo)*u
        return __dadd_rn(lo, __dmul_rn(__dsub_rn(hi, lo), sample()));
    }
    __device__ __forceinline__ void skip(int nwords) {            // consume outputs without using them
        while (nwords > 0) {
            if (pos >= MT_N) { mt_twist(s, stride); pos = 0; }
            const int take = min(nwords, MT_N - pos);
            pos += take; nwords -= take;
        }
    }
};

static __global__ void exo_seed_kernel(uint32_t* mt, int n, int stride, const uint32_t* __restrict__ seeds, int ctor_draw) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    uint32_t* s = mt + e;
    uint32_t v = seeds[e];
    s[0] = v;
    for (int i = 1; i < MT_N; i++) { v = 1812433253u * (v ^ (v >> 30)) + (uint32_t)i; s[(size_t)i * stride] = v; }
    Mt g{s, stride, MT_N};
    if (ctor_draw) (void)g.sample();                              // Exoskeleton_env.py:70
    s[(size_t)MT_N * stride] = (uint32_t)g.pos;
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

template <typename R, int NAX>
__global__ void __launch_bounds__(64)
exo_reset_kernel(EnvArrays<R> S, ChainConst<R> C, ResetConst K, StepConst<R> SK, const uint8_t* __restrict__ mask,
                 float* __restrict__ obs) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= S.n) return;
    if (mask != nullptr && mask[e] == 0) return;
    const int st = S.stride;
    const int meta = S.meta[e];
    const int mid = (meta >> 8) & 0xff;
    const int L = S.motion_len[mid];
    Mt g{S.mt + e, st, (int)S.mt[(size_t)MT_N * st + e]};

    // ---- tremor magnitude (:198-199)
    const double mag = __dadd_rn(K.amp_lo, __dmul_rn(g.sample(), __dsub_rn(K.amp_hi, K.amp_lo)));
    S.mag[e] = mag;
    // ---- generate_tremor_frequency (tremor:5-21): f1, f2, white noise, time base, the two harmonics
    const double f1 = g.uniform(K.f1_lo, K.f1_hi);
    const double f2 = g.uniform(K.f2_lo, K.f2_hi);
    const double stop = __dmul_rn((double)L, K.dt), tstep = __ddiv_rn(stop, (double)(L - 1));   // np.linspace(0, L*dt, L)
    const double w1c = __dmul_rn(__dmul_rn(2.0, 3.141592653589793), f1), w2c = __dmul_rn(__dmul_rn(2.0, 3.141592653589793), f2);
    for (int k = 0; k < L; k++) {
        S.sc_noise[(size_t)k * st + e] = __dmul_rn(g.sample(), 0.001);
        const double t = (k == L - 1) ? stop : __dmul_rn((double)k, tstep);
        S.sc_w1[(size_t)k * st + e] = sin(__dmul_rn(w1c, t));
        S.sc_w2[(size_t)k * st + e] = sin(__dmul_rn(w2c, t));
    }
    // ---- per axis: amplitudes (tremor:24-28), min/max of the raw wave (tremor:64), per-sample sign bits (tremor:70)
    double a1[7], a2[7], mn[7], mx[7];
#pragma unroll
    for (int j = 0; j < 7; j++) {
        a1[j] = pow(10.0, __ddiv_rn(g.uniform(-5.0, 0.0), 20.0));
        a2[j] = pow(10.0, __ddiv_rn(g.uniform(-20.0, -10.0), 20.0));
        if (j < NAX && ((meta >> j) & 1)) {
            uint32_t w = 0;
            for (int k = 0; k < L; k++) {
                w |= (g.next32() & 1u) << (k & 31);
                if ((k & 31) == 31 || k == L - 1) { S.sc_sign[(size_t)(j * EXO_SIGN_WORDS + (k >> 5)) * st + e] = w; w = 0; }
            }
        } else {
            g.skip(L);
        }
        mn[j] = INFINITY; mx[j] = -INFINITY;
    }
    for (int k = 0; k < L; k++) {
        const double w1 = S.sc_w1[(size_t)k * st + e], w2 = S.sc_w2[(size_t)k * st + e], nz = S.sc_noise[(size_t)k * st + e];
#pragma unroll
        for (int j = 0; j < NAX; j++) {
            const double gk = __dadd_rn(__dadd_rn(__dmul_rn(a1[j], w1), __dmul_rn(a2[j], w2)), nz);
            mn[j] = fmin(mn[j], gk); mx[j] = fmax(mx[j], gk);
        }
    }
    for (int k = 0; k < L; k++) {
        const double w1 = S.sc_w1[(size_t)k * st + e], w2 = S.sc_w2[(size_t)k * st + e], nz = S.sc_noise[(size_t)k * st + e];
#pragma unroll
        for (int j = 0; j < NAX; j++) {
            double v = 0.0;
            if ((meta >> j) & 1) {
                const double gk = __dadd_rn(__dadd_rn(__dmul_rn(a1[j], w1), __dmul_rn(a2[j], w2)), nz);
                const double jm = __dmul_rn(c_JMAX[j], mag);
                v = __dmul_rn(__dadd_rn(-1.0, __ddiv_rn(__dmul_rn(2.0, __dsub_rn(gk, mn[j])), __dsub_rn(mx[j], mn[j]))), jm);
                if (isnan(v) || isinf(v)) v = 0.0;                                                  // tremor:67
                const uint32_t w = S.sc_sign[(size_t)(j * EXO_SIGN_WORDS + (k >> 5)) * st + e];
                if (((w >> (k & 31)) & 1u) == 0) v = -v;
            }
            S.tt[((size_t)k * NAX + j) * st + e] = (R)v;
        }
    }
    // ---- add_symmetric_noise on I, D, S (:208-210): raw draws first, then (n + n^T)/2 pairwise in place
    for (int m = 0; m < 3; m++) {
        const double* M0 = (m == 0) ? c_I7 : (m == 1 ? c_D7 : c_S7);
        double* out = S.IDS + (size_t)(m * 49) * st + e;
        for (int i = 0; i < 49; i++) out[(size_t)i * st] = g.uniform(-K.noise_frac, K.noise_frac);
        for (int i = 0; i < 7; i++)
            for (int j = i; j < 7; j++) {
                const double nij = out[(size_t)(i * 7 + j) * st], nji = out[(size_t)(j * 7 + i) * st];
                const double sij = __ddiv_rn(__dadd_rn(nij, nji), 2.0), sji = __ddiv_rn(__dadd_rn(nji, nij), 2.0);
                out[(size_t)(i * 7 + j) * st] = __dadd_rn(M0[i * 7 + j], __dmul_rn(sij, M0[i * 7 + j]));
                out[(size_t)(j * 7 + i) * st] = __dadd_rn(M0[j * 7 + i], __dmul_rn(sji, M0[j * 7 + i]));
            }
    }
    // ---- actuator end shifts (sim:98-107) and force range scale (:216-217)
    for (int i = 0; i < 42; i++) S.shift[(size_t)i * st + e] = (R)g.uniform(-K.shift_range, K.shift_range);
    const double lo = __dsub_rn(1.0, K.act_range), hi = __dadd_rn(1.0, K.act_range);
    const double fsh = __dmul_rn(K.fmax_sh0, g.uniform(lo, hi));
    const double fel = __dmul_rn(K.fmax_el0, g.uniform(lo, hi));
    S.fmax[e] = (R)fsh; S.fmax[st + e] = (R)fel;
    S.mt[(size_t)MT_N * st + e] = (uint32_t)g.pos;

    // ---- counts = 2 (:220); two position reads of the unchanged link-state cache (:230-250)
    S.counts[e] = 2;
    S.ret[e] = R(0);
    R qs[5];
#pragma unroll
    for (int j = 0; j < 5; j++) qs[j] = K.lag ? S.qlag[j * st + e] : S.q[j * st + e];
    ChainPose<R> P;
    chain_fk(C, qs, P);
    R el_old[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { el_old[k] = S.elref[k * st + e]; S.elref[k * st + e] = P.elref[k]; }
    float* o = obs ? obs + (size_t)e * 80 : nullptr;
#pragma unroll
    for (int i = 0; i < 7; i++) {
        R a2[3];
        if (i < 2) affine(P.R4, P.p4, C.off_alk[i], a2);
        else affine(P.R2, P.p2, C.off_hum[i], a2);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            a2[k] += (R)S.shift[(size_t)((2 * i + 1) * 3 + k) * st + e];
            const R prev = a2[k] - (i < 2 ? el_old[k] : C.sh_ref[k]);      // stale elbow ref on the first read (:237)
            const R cur = a2[k] - (i < 2 ? P.elref[k] : C.sh_ref[k]);      // refreshed before the second (:238,249)
            S.posvec[(size_t)(i * 3 + k) * st + e] = cur;
            if (o) { o[26 + i * 3 + k] = (float)prev; o[47 + i * 3 + k] = (float)cur; }
        }
    }
    if (o) {
        for (int i = 0; i < 14; i++) o[i] = 0.0f;                          // force history zeroed (:195)
        const R nrm[4] = {R(10), R(10), R(10), R(5)};
        for (int h = 0; h < 3; h++)
            for (int j = 0; j < 4; j++) o[14 + h * 4 + j] = (float)(S.tt[((size_t)h * NAX + j) * st + e] / nrm[j]);
        for (int k = 0; k < 3; k++) {
            o[68 + k] = (float)C.sh_ref[k]; o[71 + k] = (float)P.elref[k];
            o[74 + k] = (float)C.sh_ref[k]; o[77 + k] = (float)P.elref[k];
        }
    }
    (void)SK;
}

// ---------------------------------------------------------------------------------------------------------
// Propagator: one warp per env, matrices in shared memory (fp64).
//   A = [[0, 1], [-I^-1 S, -I^-1 D]] (14x14),  B = [0; I^-1] (14x7)
//   E = e^{A h}, Phi = sum_k (A h)^k/(k+1)!  (Taylor, degree 14, h = dt/2^s with ||A h||_inf <= 1/4)
//   F = h * Phi * B, then s times: F <- (E + 1) F ; E <- E E.  G = rows 0..6 of F.
// ---------------------------------------------------------------------------------------------------------
#define PW_WARPS 4
#define PD 14

__device__ __forceinline__ void warp_matmul14(const double* A, const double* B, double* Cm, int ncol, int lane, double scale) {
    // Cm[14 x ncol] = scale * A[14x14] * B[14 x ncol]
    for (int idx = lane; idx < PD * ncol; idx += 32) {
        const int i = idx / ncol, j = idx - i * ncol;
        double acc = 0;
#pragma unroll
        for (int k = 0; k < PD; k++) acc = fma(A[i * PD + k], B[k * ncol + j], acc);
        Cm[idx] = acc * scale;
    }
}

template <typename R>
__global__ void __launch_bounds__(PW_WARPS * 32)
exo_propagator_kernel(EnvArrays<R> S, double dt, const uint8_t* __restrict__ mask) {
    __shared__ double sm[PW_WARPS][5 * PD * PD + 2 * 49 + 8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int e = blockIdx.x * PW_WARPS + warp;
    if (e >= S.n) return;
    if (mask != nullptr && mask[e] == 0) return;
    const int st = S.stride;
    double* Ah = sm[warp];
    double* E = Ah + PD * PD;
    double* Phi = E + PD * PD;
    double* P = Phi + PD * PD;
    double* T = P + PD * PD;
    double* LU = T + PD * PD;        // 49: I, factorised in place (Gauss-Jordan -> I^-1)
    double* Iinv = LU + 49;          // 49
    // ---- I^-1 by Gauss-Jordan with partial pivoting (lane-parallel over columns of [I | 1])
    for (int i = lane; i < 49; i += 32) { LU[i] = S.IDS[(size_t)i * st + e]; Iinv[i] = (i / 7 == i % 7) ? 1.0 : 0.0; }
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
        double f[7];
#pragma unroll
        for (int i = 0; i < 7; i++) f[i] = LU[i * 7 + k];
        __syncwarp();
        if (lane < 14) {
            double* M = lane < 7 ? LU : Iinv; const int j = lane % 7;
            const double pk = M[k * 7 + j];
#pragma unroll
            for (int i = 0; i < 7; i++) if (i != k) M[i * 7 + j] -= f[i] * pk;
        }
        __syncwarp();
    }
    // ---- A (unscaled) into Ah, row-sum norm
    for (int idx = lane; idx < PD * PD; idx += 32) {
        const int i = idx / PD, j = idx - i * PD;
        double v = 0;
        if (i < 7) v = (j == i + 7) ? 1.0 : 0.0;
        else {
            const double* M = S.IDS + (size_t)((j < 7 ? 98 : 49)) * st + e;   // S for the q block, D for the qdot block
            const int jj = j < 7 ? j : j - 7;
            double acc = 0;
#pragma unroll
            for (int k = 0; k < 7; k++) acc = fma(Iinv[(i - 7) * 7 + k], M[(size_t)(k * 7 + jj) * st], acc);
            v = -acc;
        }
        Ah[idx] = v;
    }
    __syncwarp();
    double nrm = 0;
    if (lane < PD) { for (int j = 0; j < PD; j++) nrm += fabs(Ah[lane * PD + j]); }
    for (int o = 16; o > 0; o >>= 1) nrm = fmax(nrm, __shfl_xor_sync(0xffffffffu, nrm, o));
    nrm *= dt;
    int s = 0;
    while (nrm > 0.25 && s < 40) { nrm *= 0.5; s++; }
    const double h = ldexp(dt, -s);
    for (int idx = lane; idx < PD * PD; idx += 32) {
        Ah[idx] *= h;
        const double id = (idx / PD == idx % PD) ? 1.0 : 0.0;
        E[idx] = id; Phi[idx] = id; P[idx] = id;
    }
    __syncwarp();
    for (int k = 1; k <= 14; k++) {
        warp_matmul14(P, Ah, T, PD, lane, 1.0 / k);      // P_k = P_{k-1} (A h) / k
        __syncwarp();
        for (int idx = lane; idx < PD * PD; idx += 32) { const double v = T[idx]; P[idx] = v; E[idx] += v; Phi[idx] += v / (k + 1); }
        __syncwarp();
    }
    // F = h * Phi * B = h * Phi[:, 7:14] * I^-1   (14 x 7), kept in P; T is scratch
    double* F = P;
    for (int idx = lane; idx < PD * 7; idx += 32) {
        const int i = idx / 7, j = idx - i * 7;
        double acc = 0;
#pragma unroll
        for (int k = 0; k < 7; k++) acc = fma(Phi[i * PD + 7 + k], Iinv[k * 7 + j], acc);
        T[idx] = acc * h;
    }
    __syncwarp();
    for (int idx = lane; idx < PD * 7; idx += 32) F[idx] = T[idx];
    __syncwarp();
    for (int it = 0; it < s; it++) {
        warp_matmul14(E, F, T, 7, lane, 1.0);            // T = E F
        __syncwarp();
        for (int idx = lane; idx < PD * 7; idx += 32) F[idx] += T[idx];
        __syncwarp();
        if (it + 1 < s) {
            warp_matmul14(E, E, T, PD, lane, 1.0);
            __syncwarp();
            for (int idx = lane; idx < PD * PD; idx += 32) E[idx] = T[idx];
            __syncwarp();
        }
    }
    for (int idx = lane; idx < 49; idx += 32) S.G[(size_t)idx * st + e] = (R)F[idx];   // rows 0..6
}

}  // namespace exo
