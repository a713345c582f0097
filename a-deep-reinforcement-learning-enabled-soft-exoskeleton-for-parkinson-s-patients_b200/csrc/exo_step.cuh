// exo_step.cuh — the fused control-step kernel: one thread per environment, one launch per env.step().
//
// Replaces, for N environments at once, Exoskeleton_env.py:368-471 (step) and what it calls:
//   transform_action :256-266 | get_actuator_positions sim:119-158 | get_force_components sim:207-298
//   get_actuator_pos_vect sim:300-336 | return_reference_dummy_pos sim:338-351 | get_torques :177-187
//   tremor lookup :403-406 | solve_diff_eq x2 :409-414 (as q = G*T, see exo_reset.cuh) | angle composition :421-424
//   set_joint_position + stepSimulation :427-433 (restated Bullet position motors, oracle/minibullet.py)
//   get_reward :341-366 | update_state_vector :487-570 | done :457
//   + the caller-side amplitude metric of Exoskeleton_agent_train.py:158-170 (forward_kinematics x3).
//
// Roofline: HBM-bound. Per env-step the kernel moves ~1.3 KB (DESIGN.md "algorithmic bytes") for ~2 kflop,
// so tensor cores are irrelevant; what matters is that every load/store below is a coalesced SoA access and
// that the row-major [n,7] / [n,80] boundary arrays go through a shared-memory transpose.
#pragma once
#include "exo_types.cuh"

namespace exo {

__device__ __forceinline__ void sincos_r(float x, float* s, float* c) { sincosf(x, s, c); }
__device__ __forceinline__ void sincos_r(double x, double* s, double* c) { sincos(x, s, c); }
__device__ __forceinline__ float exp_r(float x) { return expf(x); }
__device__ __forceinline__ double exp_r(double x) { return exp(x); }
__device__ __forceinline__ float sqrt_r(float x) { return sqrtf(x); }
__device__ __forceinline__ double sqrt_r(double x) { return sqrt(x); }
__device__ __forceinline__ float abs_r(float x) { return fabsf(x); }
__device__ __forceinline__ double abs_r(double x) { return fabs(x); }

// out = A * B * Rz(angle); A,B row-major 3x3
template <typename R>
__device__ __forceinline__ void mul_origin_rz(const R* A, const R* B, R s, R c, R* out) {
    R M[9];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) M[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        out[i * 3 + 0] = c * M[i * 3] + s * M[i * 3 + 1];
        out[i * 3 + 1] = c * M[i * 3 + 1] - s * M[i * 3];
        out[i * 3 + 2] = M[i * 3 + 2];
    }
}
template <typename R>
__device__ __forceinline__ void affine(const R* Rm, const R* p, const R* v, R* o) {
#pragma unroll
    for (int i = 0; i < 3; i++) o[i] = p[i] + (Rm[i * 3] * v[0] + Rm[i * 3 + 1] * v[1] + Rm[i * 3 + 2] * v[2]);
}

// World frames of the humerus (link 2) and forearm (link 4) and the elbow reference point for joint vector qs[5].
template <typename R>
struct ChainPose { R R2[9], p2[3], R4[9], p4[3], elref[3]; };

template <typename R>
__device__ __forceinline__ void chain_fk(const ChainConst<R>& C, const R* qs, ChainPose<R>& P) {
    R s, c, Ra[9], Rb[9], pa[3], pb[3];
    const R eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    sincos_r(qs[0], &s, &c);
    mul_origin_rz(eye, C.Rorg[0], s, c, Ra);                       // link 0
    pa[0] = C.p0[0]; pa[1] = C.p0[1]; pa[2] = C.p0[2];
    sincos_r(qs[1], &s, &c);
    affine(Ra, pa, C.porg[1], pb);
    mul_origin_rz(Ra, C.Rorg[1], s, c, Rb);                        // link 1
    sincos_r(qs[2], &s, &c);
    affine(Rb, pb, C.porg[2], P.p2);
    mul_origin_rz(Rb, C.Rorg[2], s, c, P.R2);                      // link 2 (humerus)
    sincos_r(qs[3], &s, &c);
    affine(P.R2, P.p2, C.porg[3], pa);
    mul_origin_rz(P.R2, C.Rorg[3], s, c, Ra);                      // link 3 (auxlink3)
    affine(Ra, pa, C.off_elbow, P.elref);
    sincos_r(qs[4], &s, &c);
    affine(Ra, pa, C.porg[4], P.p4);
    mul_origin_rz(Ra, C.Rorg[4], s, c, P.R4);                      // link 4 (forearm)
}

// End effector of the 7-joint Denavit-Hartenberg arm of calculate_arm_end_effector_points.py:18-50
// (alpha = +-pi/2 taken exactly, a = 0, d = (0,0,L1,0,L2,0,0)).
template <typename R>
__device__ __forceinline__ void dh_end_effector(const R* th, R L1, R L2, R* p) {
    R c0[3] = {1, 0, 0}, c1[3] = {0, 1, 0}, c2[3] = {0, 0, 1};
    p[0] = p[1] = p[2] = 0;
#pragma unroll
    for (int k = 0; k < 7; k++) {
        R s, c;
        sincos_r(th[k], &s, &c);
        const R sa = (k == 2) ? R(-1) : R(1);
        const R d = (k == 2) ? L1 : (k == 4 ? L2 : R(0));
#pragma unroll
        for (int i = 0; i < 3; i++) {
            R n0 = c * c0[i] + s * c1[i];
            R n1 = sa * c2[i];
            R n2 = sa * (s * c0[i] - c * c1[i]);
            p[i] += d * c2[i];
            c0[i] = n0; c1[i] = n1; c2[i] = n2;
        }
    }
}

template <typename R>
__device__ __forceinline__ R warp_sum(R v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

#define EXO_OBS_PAD 81

template <typename R, bool INFO, int NAX>
__global__ void __launch_bounds__(EXO_STEP_THREADS)
exo_step_kernel(EnvArrays<R> S, ChainConst<R> C, StepConst<R> K, const float* __restrict__ actions,
                float* __restrict__ obs, R* __restrict__ reward, uint8_t* __restrict__ done, InfoPtrs<R> info) {
    __shared__ float s_obs[EXO_STEP_THREADS * EXO_OBS_PAD];
    __shared__ float s_act[EXO_STEP_THREADS * 7];
    __shared__ uint8_t s_live[EXO_STEP_THREADS];

    const int tid = threadIdx.x;
    const int base = blockIdx.x * EXO_STEP_THREADS;
    const int e = base + tid;
    const int n = S.n, st = S.stride;

    // ---- stage this block's action rows (row-major [n,7]) through shared memory: coalesced
    {
        const int lim = min(EXO_STEP_THREADS, n - base) * 7;
        const float* src = actions + (size_t)base * 7;
        for (int i = tid; i < lim; i += EXO_STEP_THREADS) s_act[i] = src[i];
    }
    __syncthreads();

    bool live = false;
    if (e < n) {
        const int meta = S.meta[e];
        const int mid = (meta >> 8) & 0xff;
        const int L = S.motion_len[mid];
        const int c = S.counts[e];
        live = c < L - 1;                     // an env that already reported done idles (Exoskeleton_agent_train.py:139)
        if (!live) { done[e] = 1; reward[e] = R(0); }
        if (live) {
            float* so = s_obs + tid * EXO_OBS_PAD;
            const R* mo = S.motion + (size_t)mid * 5 * S.max_len;
            // ---- transform_action (:256-266)
            const R fm_sh = S.fmax[e], fm_el = S.fmax[st + e];
            R F[7];
#pragma unroll
            for (int i = 0; i < 7; i++) F[i] = ((R(s_act[tid * 7 + i]) + R(1)) / R(2)) * (i < 2 ? fm_el : fm_sh);
            // ---- link-state cache -> world points (getLinkState reads, sim:129-142,194-195)
            R qb[5], qs[5];
#pragma unroll
            for (int j = 0; j < 5; j++) qb[j] = S.q[j * st + e];
            if (K.lag) {
#pragma unroll
                for (int j = 0; j < 5; j++) qs[j] = S.qlag[j * st + e];
            } else {
#pragma unroll
                for (int j = 0; j < 5; j++) qs[j] = qb[j];
            }
            ChainPose<R> P;
            chain_fk(C, qs, P);
            R el_old[3];
#pragma unroll
            for (int k = 0; k < 3; k++) el_old[k] = S.elref[k * st + e];
            // previous pos_vec becomes prev_position_vectors (:385) -> obs[26:47]; new pos_vec -> obs[47:68]
#pragma unroll
            for (int i = 0; i < 21; i++) so[26 + i] = (float)S.posvec[i * st + e];
            // ---- per actuator: endpoints, force direction, pos_vec, torque (sim:207-336, env:177-187)
            R at0 = 0, at1 = 0, at2 = 0, ty0 = 0, ty1 = 0;
            const int order[7] = {0, 1, 2, 3, 4, 6, 5};          // summation order of :395-397 (3,4,5,7,6)
            const int hum_of[7] = {0, 1, 2, 3, 4, 5, 6};         // off_hum slot holding: act11,act21 (origin ends) / act32..act72 (insertion ends)
            const int fix_of[7] = {-1, -1, 0, 1, 2, 3, 4};       // fixed slot of act_i1 for the shoulder actuators
#pragma unroll
            for (int oi = 0; oi < 7; oi++) {
                const int i = order[oi];
                R a1[3], a2[3];
                if (i < 2) {
                    affine(P.R2, P.p2, C.off_hum[hum_of[i]], a1);           // act11 / act21 ride on the humerus
                    affine(P.R4, P.p4, C.off_alk[i], a2);                   // act12 / act22 ride on the forearm
                } else {
                    a1[0] = C.fixed[fix_of[i]][0]; a1[1] = C.fixed[fix_of[i]][1]; a1[2] = C.fixed[fix_of[i]][2];
                    affine(P.R2, P.p2, C.off_hum[hum_of[i]], a2);
                }
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    a1[k] += S.shift[((2 * i) * 3 + k) * st + e];
                    a2[k] += S.shift[((2 * i + 1) * 3 + k) * st + e];
                }
                R dx, dy, dz;
                if (sizeof(R) == 8) {   // the reference's +5 shift (sim:212) only matters at fp64 resolution
                    dx = (a2[0] + R(5)) - (a1[0] + R(5)); dy = (a2[1] + R(5)) - (a1[1] + R(5)); dz = (a2[2] + R(5)) - (a1[2] + R(5));
                } else { dx = a2[0] - a1[0]; dy = a2[1] - a1[1]; dz = a2[2] - a1[2]; }
                // cos(atan2(dy,dx)), cos(atan2(dx,dy)), cos(atan2(dz,dx))  (not a unit vector; reproduced as is)
                const R rxy = sqrt_r(dx * dx + dy * dy), rxz = sqrt_r(dx * dx + dz * dz);
                const R fx = F[i] * (rxy > R(0) ? dx / rxy : R(1));
                const R fy = F[i] * (rxy > R(0) ? dy / rxy : R(1));
                const R fz = F[i] * (rxz > R(0) ? dx / rxz : R(1));
                const R* stale = (i < 2) ? el_old : C.sh_ref;               // refs cached by the previous read (sim:300-336)
                const R* fresh = (i < 2) ? P.elref : C.sh_ref;
                R pv[3], r[3];
#pragma unroll
                for (int k = 0; k < 3; k++) { pv[k] = a2[k] - stale[k]; r[k] = fresh[k] - a2[k]; }
#pragma unroll
                for (int k = 0; k < 3; k++) { S.posvec[(i * 3 + k) * st + e] = pv[k]; so[47 + i * 3 + k] = (float)pv[k]; }
                const R tx = fy * r[2] - fz * r[1], ty = fz * r[0] - fx * r[2], tz = fx * r[1] - fy * r[0];   // cross(F, r) :179
                if (i == 0) ty0 = ty;
                else if (i == 1) ty1 = ty;
                else { at0 += ty; at1 += tx; at2 += tz; }
            }
            const R at[4] = {at0, at1, at2, abs_r(ty0) - abs_r(ty1)};      // :394-400
            // ---- tremor lookup (:403-406) and the rows the observation needs (:502-513)
            const R* tt = S.tt;
            R ttc[7], tv[7];
#pragma unroll
            for (int j = 0; j < 7; j++) {
                ttc[j] = (j < NAX) ? tt[((size_t)c * NAX + j) * st + e] : R(0);
                tv[j] = ttc[j] + (j < 4 ? at[j] : R(0));
            }
            const R nrm[4] = {R(10), R(10), R(10), R(5)};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                so[14 + j] = (float)(tt[((size_t)(c - 1) * NAX + j) * st + e] / nrm[j]);
                so[18 + j] = (float)(ttc[j] / nrm[j]);
                so[22 + j] = (float)(tt[((size_t)(c + 1) * NAX + j) * st + e] / nrm[j]);
            }
            // ---- joint model: q(dt) = G * T from rest (:409-418); rows 0..3 drive the motors
            const R r2d = R(57.29577951308232087679815481410517), d2r = R(0.017453292519943295769236907684886);
            R qa[7], qt[7];
            constexpr int ROWS = INFO ? 7 : 4;
#pragma unroll
            for (int i = 0; i < ROWS; i++) {
                R acc = 0, acc2 = 0;
#pragma unroll
                for (int j = 0; j < 7; j++) {
                    if (j < 4 || j < NAX) {
                        const R g = S.G[(i * 7 + j) * st + e];
                        acc += g * tv[j];
                        if (INFO) acc2 += g * ttc[j];
                    }
                }
                qa[i] = acc * r2d;
                if (INFO) qt[i] = acc2 * r2d;
            }
            // ---- angle composition + position motors + one Bullet step (:421-433, oracle/minibullet.py)
            const R tgt[5] = {(mo[4 * S.max_len + c] + qa[2]) * d2r, (mo[3 * S.max_len + c] + qa[0]) * d2r,
                              (mo[2 * S.max_len + c] + qa[1]) * d2r, (mo[0 * S.max_len + c] + qa[3]) * d2r,
                              mo[1 * S.max_len + c] * d2r};
            int crossed = 0;
#pragma unroll
            for (int j = 0; j < 5; j++) {
                R qn = qb[j] + R(0.025) * ((R(0.1) * (tgt[j] - qb[j])) / R(0.025));   // q + dt*v*, v* = kp*(target-q)/dt
                if (qn < C.lower[j] || qn > C.upper[j]) {
                    crossed++;
                    if (K.limits) qn = qn < C.lower[j] ? C.lower[j] : C.upper[j];
                }
                S.q[j * st + e] = qn;
                if (K.lag) S.qlag[j * st + e] = qb[j];
            }
#pragma unroll
            for (int k = 0; k < 3; k++) S.elref[k * st + e] = P.elref[k];
            // ---- reward (:268-366)
            const R eps = R(1e-10);
            const R m = fm_el + fm_sh;
            int nax = 0;
#pragma unroll
            for (int j = 0; j < 7; j++) nax += (meta >> j) & 1;
            const R nf = (R)nax;
            R sum_unw = 0, sum_tor = 0, sumF = 0, smooth = 0;
            int axis_cnt = 0, arm_red = 0;
#pragma unroll
            for (int j = 0; j < 7; j++) {
                const R a = abs_r(tv[j]), b = abs_r(ttc[j]);
                if (j < 4) {
                    if ((meta >> j) & 1) sum_tor += (a - b) / b + R(1);
                    else sum_unw += a;
                }
                const R red = (a - b) / (b + eps) * R(100);
                if (red < R(0)) { axis_cnt++; if (j < 4) arm_red = 1; }      // nan/inf -> 0 -> not counted (:302-303)
            }
            R fp1[7];
#pragma unroll
            for (int i = 0; i < 7; i++) {
                fp1[i] = S.fprev[i * st + e];
                const R fp2 = S.fprev2[i * st + e];
                const R d = F[i] - R(2) * fp1[i] + fp2;
                smooth += d * d;
                sumF += F[i];
                S.fprev2[i * st + e] = fp1[i];
                S.fprev[i * st + e] = F[i];
            }
            const R r_unw = exp_r(-(sum_unw / (m / R(4) / nf)) + eps) * R(0.5);
            const R r_tor = exp_r((-sum_tor + eps) / nf) * R(0.9);
            const R r_axis = (R)axis_cnt * R(0.5);
            const R r_ctl = exp_r(-(sumF / (m / R(2))) + eps) * R(0.05);
            const R r_smo = R(0.05) * exp_r(-((smooth / R(7)) / (m / R(4))) + eps);
            const R max_reward = nf * R(0.5) + R(0.9) + R(0.05) + R(0.05) + R(0.5);
            const R rew = (r_axis + r_tor + r_smo + r_ctl + r_unw) / max_reward;
            reward[e] = rew;
            // ---- counts, observation head, done (:448-457)
            const int k1 = c + 1;
            S.counts[e] = k1;
#pragma unroll
            for (int i = 0; i < 7; i++) {
                const R f0 = i < 2 ? K.fmax_el0 : K.fmax_sh0;
                so[i] = (c >= 3) ? (float)(fp1[i] / f0) : 0.0f;       // forces[k-2]; the history is zeroed at reset (:195)
                so[7 + i] = (float)(F[i] / f0);                       // forces[k-1]
            }
#pragma unroll
            for (int k = 0; k < 3; k++) {
                so[68 + k] = (float)C.sh_ref[k]; so[71 + k] = (float)el_old[k];     // ref_prev (:383)
                so[74 + k] = (float)C.sh_ref[k]; so[77 + k] = (float)P.elref[k];    // ref (:387)
            }
            const bool dn = k1 >= L - 1;
            done[e] = dn ? 1 : 0;
            // ---- episode statistics (device-side; all-reduced over NCCL by the host)
            const R ret = S.ret[e] + rew;
            S.ret[e] = dn ? R(0) : ret;
            R amp = 0;
            if (INFO || K.metrics) {
                if (INFO) {
                    // Exoskeleton_agent_train.py:158-170 with return_original_joint_angles at the incremented counts
                    R b[7] = {mo[2 * S.max_len + k1] * d2r, mo[3 * S.max_len + k1] * d2r, mo[4 * S.max_len + k1] * d2r,
                              mo[0 * S.max_len + k1] * d2r, mo[1 * S.max_len + k1] * d2r, R(0), R(0)};
                    R th[7], p0[3], p1[3], p2[3];
                    dh_end_effector(b, K.L1, K.L2, p0);
#pragma unroll
                    for (int i = 0; i < 7; i++) th[i] = qa[i] * d2r + b[i];
                    dh_end_effector(th, K.L1, K.L2, p1);
#pragma unroll
                    for (int i = 0; i < 7; i++) th[i] = qt[i] * d2r + b[i];
                    dh_end_effector(th, K.L1, K.L2, p2);
                    const R ds = sqrt_r((p1[0] - p0[0]) * (p1[0] - p0[0]) + (p1[1] - p0[1]) * (p1[1] - p0[1]) + (p1[2] - p0[2]) * (p1[2] - p0[2]));
                    const R du = sqrt_r((p2[0] - p0[0]) * (p2[0] - p0[0]) + (p2[1] - p0[1]) * (p2[1] - p0[1]) + (p2[2] - p0[2]) * (p2[2] - p0[2]));
                    amp = ((ds - du) / du) * R(100);
                }
            }
            if (INFO) {
#pragma unroll
                for (int j = 0; j < 7; j++) {
                    if (info.actuator_torques) info.actuator_torques[(size_t)e * 7 + j] = j < 4 ? at[j] : R(0);
                    if (info.torque_val) info.torque_val[(size_t)e * 7 + j] = tv[j];
                    if (info.ampl_val) info.ampl_val[(size_t)e * 7 + j] = qa[j];
                    if (info.tremor_torque_val) info.tremor_torque_val[(size_t)e * 7 + j] = ttc[j];
                    if (info.tremor_ampl_val) info.tremor_ampl_val[(size_t)e * 7 + j] = qt[j];
                }
                if (info.reward_parts) {
                    R* rp = info.reward_parts + (size_t)e * 5;
                    rp[0] = r_unw; rp[1] = r_tor; rp[2] = r_axis; rp[3] = r_ctl; rp[4] = r_smo;
                }
                if (info.amp_metric) info.amp_metric[e] = amp;
            }
            // warp-aggregated statistics: one RED per warp and counter
            {
                const unsigned mask = __activemask();
                const int lane = tid & 31;
                const int leader = __ffs(mask) - 1;
                double v_ep = dn ? 1.0 : 0.0, v_ret = dn ? (double)ret : 0.0, v_len = dn ? (double)(k1 - 2) : 0.0;
                double v_cross = (double)crossed, v_red = (double)arm_red;
                double v_amp = (INFO && amp < R(0)) ? (double)amp : 0.0, v_ampn = (INFO && amp < R(0)) ? 1.0 : 0.0;
                double v_steps = 1.0;
                for (int o = 16; o > 0; o >>= 1) {
                    v_ep += __shfl_xor_sync(mask, v_ep, o); v_ret += __shfl_xor_sync(mask, v_ret, o);
                    v_len += __shfl_xor_sync(mask, v_len, o); v_cross += __shfl_xor_sync(mask, v_cross, o);
                    v_red += __shfl_xor_sync(mask, v_red, o); v_steps += __shfl_xor_sync(mask, v_steps, o);
                    if (INFO) { v_amp += __shfl_xor_sync(mask, v_amp, o); v_ampn += __shfl_xor_sync(mask, v_ampn, o); }
                }
                if (mask == 0xffffffffu && lane == leader) {
                    if (v_ep != 0.0) { atomicAdd(S.stats + 0, v_ep); atomicAdd(S.stats + 1, v_ret); atomicAdd(S.stats + 2, v_len); }
                    atomicAdd(S.stats + 3, v_steps);
                    if (v_cross != 0.0) atomicAdd(S.stats + 4, v_cross);
                    if (v_red != 0.0) atomicAdd(S.stats + 5, v_red);
                    if (INFO && v_ampn != 0.0) { atomicAdd(S.stats + 6, v_amp); atomicAdd(S.stats + 7, v_ampn); }
                } else if (mask != 0xffffffffu) {
                    // partially active warp (tail block / mixed idle): xor-shuffles over a partial mask are not a
                    // reduction, so every active lane contributes its own values
                    if (dn) { atomicAdd(S.stats + 0, 1.0); atomicAdd(S.stats + 1, (double)ret); atomicAdd(S.stats + 2, (double)(k1 - 2)); }
                    atomicAdd(S.stats + 3, 1.0);
                    if (crossed) atomicAdd(S.stats + 4, (double)crossed);
                    if (arm_red) atomicAdd(S.stats + 5, 1.0);
                    if (INFO && amp < R(0)) { atomicAdd(S.stats + 6, (double)amp); atomicAdd(S.stats + 7, 1.0); }
                }
            }
        }
    }
    s_live[tid] = live ? 1 : 0;
    __syncthreads();
    // ---- observation tile -> row-major [n,80] with 128-bit coalesced stores
    if (obs != nullptr) {
        const int rows = min(EXO_STEP_THREADS, n - base);
        float4* dst = reinterpret_cast<float4*>(obs + (size_t)base * 80);
        for (int v = tid; v < rows * 20; v += EXO_STEP_THREADS) {
            const int row = v / 20, col = (v - row * 20) * 4;
            if (s_live[row]) {
                const float* sp = s_obs + row * EXO_OBS_PAD + col;
                dst[v] = make_float4(sp[0], sp[1], sp[2], sp[3]);
            }
        }
    }
}

}  // namespace exo
