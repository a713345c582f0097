// exo_motor.cuh — the general Bullet multibody step for the exo_v3.urdf tree, one thread per environment
// (ExoConfig.motor_model = EXO_MOTOR_FULL). Replaces `client.stepSimulation()` (Exoskeleton_env.py:433) after
// `setJointMotorControlArray(POSITION_CONTROL)` (Exoskeleton_sim_pybullet.py:109-117) WITHOUT assuming that the
// constraint solver converges: forward dynamics of the tree (gravity, Bullet's per-link velocity damping, Coriolis /
// centrifugal / gyroscopic terms), one motor row per joint with Bullet's impulse caps, joint-limit rows while a bound
// is violated, `iterations` sweeps of projected Gauss-Seidel in Bullet's row order with its early-exit test, then
// semi-implicit Euler. The algorithm is restated in oracle/multibody.py (fp64 numpy; its header lists what each stage
// follows in Bullet); tests/test_gpu_parity.py compares the two. The default, EXO_MOTOR_IDEAL, is the fixed point of
// this iteration (q+ = q + kp (target - q)), evaluated inside the step kernel.
//
// Degrees of freedom: 5 revolute arm joints + 9 prismatic dummies carried by the humerus (7) and forearm (2). The 5
// dummies on the fixed base are decoupled 1-DoF systems (M row = 1 kg on the diagonal): their default velocity motor
// cancels the gravity impulse (0.245 N s < the 1 N s cap) exactly in its first sweep, so they never move and are not
// simulated. Everything is evaluated in the world frame.
//
// Cost: ~12 kflop + up to 50 x 14 x 15 FMA per env-step, fp32 or fp64 like the rest of the handle; the symmetric 14 x 14
// inverse mass matrix lives in shared memory ([entry][thread], 105 entries, compile-time offsets). Compute-bound, unlike
// the HBM-bound default step.
#pragma once
#include "exo_step.cuh"

namespace exo {

template <typename R> __device__ __forceinline__ void cross3(const R* a, const R* b, R* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
template <typename R> __device__ __forceinline__ R dot3(const R* a, const R* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename R> __device__ __forceinline__ void matvec3(const R* M, const R* v, R* o) {
#pragma unroll
    for (int i = 0; i < 3; i++) o[i] = M[i * 3] * v[0] + M[i * 3 + 1] * v[1] + M[i * 3 + 2] * v[2];
}
// world inertia tensor Rm * I6 * Rm^T (I6 = xx,xy,xz,yy,yz,zz in the body frame) -> full 3x3
template <typename R> __device__ __forceinline__ void rotate_inertia(const R* Rm, const R* I6, R* out) {
    const R I[9] = {I6[0], I6[1], I6[2], I6[1], I6[3], I6[4], I6[2], I6[4], I6[5]};
    R T[9];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) T[i * 3 + j] = Rm[i * 3] * I[j] + Rm[i * 3 + 1] * I[3 + j] + Rm[i * 3 + 2] * I[6 + j];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) out[i * 3 + j] = T[i * 3] * Rm[j * 3] + T[i * 3 + 1] * Rm[j * 3 + 1] + T[i * 3 + 2] * Rm[j * 3 + 2];
}

#define EXO_MOTOR_THREADS 64
#ifndef EXO_MOTOR_BLOCKS
#define EXO_MOTOR_BLOCKS 6      // resident 64-thread blocks per SM the fp32 instantiation is compiled for (<= 170 registers)
#endif

#define EXO_NTRI (EXO_NDOF * (EXO_NDOF + 1) / 2)
#define EXO_NCONS 19             // 14 motor rows + the 5 revolute joints' limit constraints: the static solver path's list length

// The common case - no joint of any env of the warp at a bound (84 % of warps in the bench population) - leaves only the 14
// motor rows, whose order is the solver-slot order by construction: every index below is a compile-time constant, the
// velocity change is kept per slot, and a row costs 14 loads + FMAs. `big` collects "some row moved more than the exit
// threshold": |dl| / dinv > sqrt(threshold), tested as |dl| > lim[r] = sqrt(threshold) * dinv[r] (fp32 handles; the fp64
// instantiation keeps the CPU model's exact form res * res > threshold so that sweep counts agree bit for bit).
__host__ __device__ constexpr int tri_c(int a, int b) { return a >= b ? a * (a + 1) / 2 + b : b * (b + 1) / 2 + a; }   // == DynConst::tri

template <typename R, bool FWD>
__device__ __forceinline__ void pgs_sweep_motors(const R* sM, const DynConst<R>& D, const R* dinv, const R* rhs, const R* lim,
                                                 R* lam, R* dv, R& worst, bool& big) {
    constexpr int NT = EXO_MOTOR_THREADS;
#pragma unroll
    for (int k = 0; k < EXO_NDOF; k++) {
        const int r = FWD ? k : EXO_NDOF - 1 - k;
        const R cap = D.slot_dof[r] < 5 ? D.cap_pos : D.cap_vel;
        R dl = rhs[r] - dv[r] * dinv[r];
        const R sum = lam[r] + dl;
        if (sum < -cap) { dl = -cap - lam[r]; lam[r] = -cap; }
        else if (sum > cap) { dl = cap - lam[r]; lam[r] = cap; }
        else lam[r] = sum;
#pragma unroll
        for (int i = 0; i < EXO_NDOF; i++) dv[i] += sM[tri_c(i, r) * NT] * dl;       // compile-time offsets: LDS + FFMA per element
        if (sizeof(R) == 8) { const R res = dl / dinv[r]; worst = fmax(worst, res * res); }
        else big |= abs_r(dl) > lim[r];
    }
}

// One projected Gauss-Seidel sweep over Bullet's constraint list (19 entries, forward or backward). Entry ci is a motor row
// (always present) or a joint-limit constraint (a row only in the threads where that joint is at / over a bound: bit ci of
// `act`; bit ci of `neg` = the upper bound). EVERYTHING is indexed by the compile-time entry index ci: rhs, 1/diag, impulse and
// dvc[ci] = the velocity change so far of the joint entry ci acts on (a limit entry shares its joint with a motor entry and
// carries its own copy). A row's impulse updates all 19 copies through D.etri[c2][ci] = the M^-1 entry (joint of c2, joint
// of ci) - 19 loads + FMAs with constant-bank offsets, no run-time register index. (The first version kept dv per joint and
// fetched a row's own dv through a 14-way switch: 28 % of the kernel's instructions, r02B.)
template <typename R, bool FWD>
__device__ __forceinline__ void pgs_sweep(const R* sM, const DynConst<R>& D, unsigned act, unsigned neg, const R* dinv,
                                          const R* rhs, R* lam, R* dvc, R& worst) {
    constexpr int NT = EXO_MOTOR_THREADS;
#pragma unroll
    for (int k = 0; k < EXO_NCONS; k++) {
        const int ci = FWD ? k : EXO_NCONS - 1 - k;
        const int cc = D.cons[ci];                                  // warp-uniform
        R imp, dl;
        if (cc >= 0) {
            const R cap = cc < 5 ? D.cap_pos : D.cap_vel;
            dl = rhs[ci] - dvc[ci] * dinv[ci];
            const R sum = lam[ci] + dl;
            if (sum < -cap) { dl = -cap - lam[ci]; lam[ci] = -cap; }
            else if (sum > cap) { dl = cap - lam[ci]; lam[ci] = cap; }
            else lam[ci] = sum;
            imp = dl;
        } else {
            if (!((act >> ci) & 1u)) continue;                       // divergent only in warps where a joint is at a limit
            const R sg = ((neg >> ci) & 1u) ? R(-1) : R(1);
            dl = rhs[ci] - sg * dvc[ci] * dinv[ci];
            const R sum = lam[ci] + dl;
            if (sum < R(0)) { dl = -lam[ci]; lam[ci] = 0; }
            else if (sum > D.limit_cap) { dl = D.limit_cap - lam[ci]; lam[ci] = D.limit_cap; }
            else lam[ci] = sum;
            imp = sg * dl;
        }
#pragma unroll
        for (int c2 = 0; c2 < EXO_NCONS; c2++) dvc[c2] += sM[D.etri[c2][ci] * NT] * imp;
        const R res = div_r(dl, dinv[ci]);                        // fp64: IEEE division as in the CPU model; fp32: fast division
        worst = fmax(worst, res * res);
    }
}

template <typename R>
__global__ void __launch_bounds__(EXO_MOTOR_THREADS, sizeof(R) == 4 ? EXO_MOTOR_BLOCKS : 2)
exo_motor_kernel(EnvArrays<R> S, ChainConst<R> C, DynConst<R> D, int lag, int env0, int env1) {
    constexpr int NT = EXO_MOTOR_THREADS;
    extern __shared__ __align__(16) unsigned char motor_smem[];
    // M^-1 is symmetric: its 105 distinct entries in solver-slot order, entry D.tri[a][b] of thread t at [entry][t]. (The
    // full 14 x 14 image, 54 KB per 64-thread block, held an SM to 8 warps; 30 KB lets all 14 warps of a 65 536-env launch be
    // resident at once, and the solver is a dependent chain that only other warps can hide.)
    R* sM = reinterpret_cast<R*>(motor_smem) + threadIdx.x;                       // [105][NT]
    R* sV = reinterpret_cast<R*>(motor_smem) + EXO_NTRI * NT + threadIdx.x;        // [14][NT]: reorder scratch
    const int e = env0 + blockIdx.x * blockDim.x + threadIdx.x;
    const int st = S.stride;
    int sweeps_done = 0, did = 0;
    if (e < env1) {
        R tg[8];
        ld4(S.tgt + e, tg); ld4(S.tgt + (size_t)st + e, tg + 4);
        if (tg[5] != R(0)) {
            did = 1;
            R q[5], dy[4 * NV_DYN];
#pragma unroll
            for (int j = 0; j < 5; j++) q[j] = *exo_word(S.state, W_Q + j, e, st);
#pragma unroll
            for (int v = 0; v < NV_DYN; v++) ld4(S.dyn + (size_t)v * st + e, dy + 4 * v);
            R* qd = dy + W_QD;
            R* qp = dy + W_QP;
            // ---- 1. kinematics of the 5 revolute links
            R Rl[5][9], pl[5][3], ax[5][3];
            {
                const R eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
                R sn, cs;
                sincos_r(q[0], &sn, &cs);
                mul_origin_rz(eye, C.Rorg[0], sn, cs, Rl[0]);
                pl[0][0] = C.p0[0]; pl[0][1] = C.p0[1]; pl[0][2] = C.p0[2];
#pragma unroll
                for (int i = 1; i < 5; i++) {
                    sincos_r(q[i], &sn, &cs);
                    affine(Rl[i - 1], pl[i - 1], C.porg[i], pl[i]);
                    mul_origin_rz(Rl[i - 1], C.Rorg[i], sn, cs, Rl[i]);
                }
#pragma unroll
                for (int i = 0; i < 5; i++) { ax[i][0] = Rl[i][2]; ax[i][1] = Rl[i][5]; ax[i][2] = Rl[i][8]; }   // joint axis = local z
            }
            // ---- 2. velocities / zero-qdd accelerations of the links (recursive Newton-Euler, forward pass)
            R w[5][3], al[5][3], vo[5][3], ao[5][3];
            {
                R pw[3] = {0, 0, 0}, pal[3] = {0, 0, 0}, pv[3] = {0, 0, 0}, pa[3] = {0, 0, 0};
#pragma unroll
                for (int i = 0; i < 5; i++) {
                    R r[3] = {0, 0, 0};
                    if (i > 0) { r[0] = pl[i][0] - pl[i - 1][0]; r[1] = pl[i][1] - pl[i - 1][1]; r[2] = pl[i][2] - pl[i - 1][2]; }
                    R wr[3], wwr[3], ar[3], wa[3];
                    cross3(pw, r, wr); cross3(pw, wr, wwr); cross3(pal, r, ar); cross3(pw, ax[i], wa);
#pragma unroll
                    for (int k = 0; k < 3; k++) {
                        vo[i][k] = pv[k] + wr[k];
                        ao[i][k] = pa[k] + ar[k] + wwr[k];
                        w[i][k] = pw[k] + ax[i][k] * qd[i];
                        al[i][k] = pal[k] + wa[k] * qd[i];
                    }
#pragma unroll
                    for (int k = 0; k < 3; k++) { pw[k] = w[i][k]; pal[k] = al[i][k]; pv[k] = vo[i][k]; pa[k] = ao[i][k]; }
                }
            }
            // ---- 3. per body: inertial force/torque (gravity, damping, gyroscopic) and the composite inertias of M
            R fl[5][3], nl[5][3];                       // force / moment about the link origin, accumulated over the subtree
            R cm[5], ch[5][3], cI[5][6];                // composite mass, first moment, inertia about the world origin (per link, own bodies)
            R bias[EXO_NDOF];
            R B[5][EXO_NPRIS];                          // M[rev k][dummy d]
#pragma unroll
            for (int i = 0; i < 5; i++) {
#pragma unroll
                for (int k = 0; k < 3; k++) { fl[i][k] = 0; nl[i][k] = 0; ch[i][k] = 0; }
#pragma unroll
                for (int k = 0; k < 6; k++) cI[i][k] = 0;
                cm[i] = 0;
            }
            auto body = [&](int carrier, R m, const R* c, const R* Iw, const R* vc, const R* ac, R* F, R* N) {
                // Newton-Euler of one rigid body riding on `carrier` (angular state of the carrier)
                R Iwv[3], Ial[3], gy[3];
                matvec3(Iw, w[carrier], Iwv); matvec3(Iw, al[carrier], Ial); cross3(w[carrier], Iwv, gy);
                const R nv = sqrt_r(dot3(vc, vc)), nw = sqrt_r(dot3(w[carrier], w[carrier]));
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    F[k] = m * (ac[k] - D.g[k]) + m * vc[k] * (D.k_lin + D.k_lin * nv);
                    N[k] = Ial[k] + gy[k] + Iwv[k] * (D.k_ang + D.k_ang * nw);
                }
                // composite inertia of the carrier's own bodies about the carrier's joint origin (no large offsets: fp32-safe)
                const R rr[3] = {c[0] - pl[carrier][0], c[1] - pl[carrier][1], c[2] - pl[carrier][2]};
                const R r2 = dot3(rr, rr);
                cm[carrier] += m;
#pragma unroll
                for (int k = 0; k < 3; k++) ch[carrier][k] += m * rr[k];
                cI[carrier][0] += Iw[0] + m * (r2 - rr[0] * rr[0]); cI[carrier][1] += Iw[1] - m * rr[0] * rr[1];
                cI[carrier][2] += Iw[2] - m * rr[0] * rr[2];        cI[carrier][3] += Iw[4] + m * (r2 - rr[1] * rr[1]);
                cI[carrier][4] += Iw[5] - m * rr[1] * rr[2];        cI[carrier][5] += Iw[8] + m * (r2 - rr[2] * rr[2]);
            };
#pragma unroll
            for (int i = 0; i < 5; i++) {               // the links themselves
                R rc[3], c[3], Iw[9], t1[3], t2[3], t3[3], vc[3], ac[3], F[3], N[3], rxF[3];
                matvec3(Rl[i], D.com_link[i], rc);
#pragma unroll
                for (int k = 0; k < 3; k++) c[k] = pl[i][k] + rc[k];
                rotate_inertia(Rl[i], D.I_link[i], Iw);
                cross3(w[i], rc, t1); cross3(w[i], t1, t2); cross3(al[i], rc, t3);
#pragma unroll
                for (int k = 0; k < 3; k++) { vc[k] = vo[i][k] + t1[k]; ac[k] = ao[i][k] + t3[k] + t2[k]; }
                body(i, D.m_link[i], c, Iw, vc, ac, F, N);
                cross3(rc, F, rxF);
#pragma unroll
                for (int k = 0; k < 3; k++) { fl[i][k] += F[k]; nl[i][k] += N[k] + rxF[k]; }
            }
            R ud[EXO_NPRIS][3], cd[EXO_NPRIS][3];
#pragma unroll
            for (int d = 0; d < EXO_NPRIS; d++) {        // the dummies: prismatic leaves on link 2 (humerus) or 4 (forearm)
                const int cr = d < 7 ? 2 : 4;
                const R* off = d < 7 ? C.off_hum[d] : C.off_alk[d - 7];
                const R* pax = d < 7 ? C.pax_hum[d] : C.pax_alk[d - 7];
                R loc[3], r[3], Iw[9], t1[3], t2[3], t3[3], t4[3], vc[3], ac[3], F[3], N[3], rxF[3];
#pragma unroll
                for (int k = 0; k < 3; k++) loc[k] = off[k] + pax[k] * qp[d];
                matvec3(Rl[cr], loc, r); matvec3(Rl[cr], pax, ud[d]);
#pragma unroll
                for (int k = 0; k < 3; k++) cd[d][k] = pl[cr][k] + r[k];
                rotate_inertia(Rl[cr], D.I_d[d], Iw);
                cross3(w[cr], r, t1); cross3(w[cr], t1, t2); cross3(al[cr], r, t3); cross3(w[cr], ud[d], t4);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    vc[k] = vo[cr][k] + t1[k] + ud[d][k] * qd[5 + d];
                    ac[k] = ao[cr][k] + t3[k] + t2[k] + R(2) * t4[k] * qd[5 + d];
                }
                body(cr, D.m_d[d], cd[d], Iw, vc, ac, F, N);
                bias[5 + d] = dot3(ud[d], F);
                cross3(r, F, rxF);
#pragma unroll
                for (int k = 0; k < 3; k++) { fl[cr][k] += F[k]; nl[cr][k] += N[k] + rxF[k]; }
#pragma unroll
                for (int k = 0; k < 5; k++) {
                    R rel[3] = {cd[d][0] - pl[k][0], cd[d][1] - pl[k][1], cd[d][2] - pl[k][2]}, x[3];
                    cross3(rel, ud[d], x);
                    B[k][d] = (k <= cr) ? D.m_d[d] * dot3(ax[k], x) : R(0);
                }
            }
            // backward passes: joint torques of the bias, composite inertias -> revolute block of M
            R Mrr[5][5];
#pragma unroll
            for (int i = 4; i >= 0; i--) {
                bias[i] = dot3(ax[i], nl[i]);
                if (i > 0) {
                    R r[3] = {pl[i][0] - pl[i - 1][0], pl[i][1] - pl[i - 1][1], pl[i][2] - pl[i - 1][2]}, rxf[3];
                    cross3(r, fl[i], rxf);
#pragma unroll
                    for (int k = 0; k < 3; k++) { fl[i - 1][k] += fl[i][k]; nl[i - 1][k] += nl[i][k] + rxf[k]; }
                }
                // unit velocity of joint i moves its subtree (composite about p_i): momentum P, angular momentum about p_i
                R P[3];
                cross3(ax[i], ch[i], P);
                const R Lp[3] = {cI[i][0] * ax[i][0] + cI[i][1] * ax[i][1] + cI[i][2] * ax[i][2],
                                 cI[i][1] * ax[i][0] + cI[i][3] * ax[i][1] + cI[i][4] * ax[i][2],
                                 cI[i][2] * ax[i][0] + cI[i][4] * ax[i][1] + cI[i][5] * ax[i][2]};
#pragma unroll
                for (int k = 0; k <= i; k++) {
                    const R rel[3] = {pl[i][0] - pl[k][0], pl[i][1] - pl[k][1], pl[i][2] - pl[k][2]};
                    R rxP[3];
                    cross3(rel, P, rxP);
                    const R v = ax[k][0] * (Lp[0] + rxP[0]) + ax[k][1] * (Lp[1] + rxP[1]) + ax[k][2] * (Lp[2] + rxP[2]);
                    Mrr[k][i] = v; Mrr[i][k] = v;
                }
                if (i > 0) {                            // shift the composite of link i to the origin of link i-1 and add it there
                    const R r[3] = {pl[i][0] - pl[i - 1][0], pl[i][1] - pl[i - 1][1], pl[i][2] - pl[i - 1][2]};
                    const R* h = ch[i];
                    const R r2 = dot3(r, r), hr = dot3(h, r), mi = cm[i];
                    cI[i - 1][0] += cI[i][0] + mi * (r2 - r[0] * r[0]) + R(2) * hr - R(2) * h[0] * r[0];
                    cI[i - 1][1] += cI[i][1] - mi * r[0] * r[1] - h[0] * r[1] - r[0] * h[1];
                    cI[i - 1][2] += cI[i][2] - mi * r[0] * r[2] - h[0] * r[2] - r[0] * h[2];
                    cI[i - 1][3] += cI[i][3] + mi * (r2 - r[1] * r[1]) + R(2) * hr - R(2) * h[1] * r[1];
                    cI[i - 1][4] += cI[i][4] - mi * r[1] * r[2] - h[1] * r[2] - r[1] * h[2];
                    cI[i - 1][5] += cI[i][5] + mi * (r2 - r[2] * r[2]) + R(2) * hr - R(2) * h[2] * r[2];
#pragma unroll
                    for (int k = 0; k < 3; k++) ch[i - 1][k] += h[k] + mi * r[k];
                    cm[i - 1] += mi;
                }
            }
            // ---- 4. M^-1 through the Schur complement of the (diagonal) dummy block
            R Sc[5][5];
#pragma unroll
            for (int k = 0; k < 5; k++)
#pragma unroll
                for (int l = 0; l < 5; l++) {
                    R acc = Mrr[k][l];
#pragma unroll
                    for (int d = 0; d < EXO_NPRIS; d++) acc -= B[k][d] * B[l][d] / D.m_d[d];
                    Sc[k][l] = acc;
                }
            R Lc[5][5], Li[5][5], Sinv[5][5];          // Cholesky S = L L^T, L^-1, S^-1 = L^-T L^-1
#pragma unroll
            for (int i = 0; i < 5; i++)
#pragma unroll
                for (int j = 0; j < 5; j++) { Lc[i][j] = 0; Li[i][j] = 0; }
#pragma unroll
            for (int j = 0; j < 5; j++) {
                R dsum = Sc[j][j];
#pragma unroll
                for (int k = 0; k < j; k++) dsum -= Lc[j][k] * Lc[j][k];
                Lc[j][j] = sqrt_r(dsum);
#pragma unroll
                for (int i = j + 1; i < 5; i++) {
                    R acc = Sc[i][j];
#pragma unroll
                    for (int k = 0; k < j; k++) acc -= Lc[i][k] * Lc[j][k];
                    Lc[i][j] = acc / Lc[j][j];
                }
            }
#pragma unroll
            for (int j = 0; j < 5; j++) {
                Li[j][j] = R(1) / Lc[j][j];
#pragma unroll
                for (int i = j + 1; i < 5; i++) {
                    R acc = 0;
#pragma unroll
                    for (int k = j; k < i; k++) acc -= Lc[i][k] * Li[k][j];
                    Li[i][j] = acc / Lc[i][i];
                }
            }
#pragma unroll
            for (int i = 0; i < 5; i++)
#pragma unroll
                for (int j = 0; j < 5; j++) {
                    R acc = 0;
#pragma unroll
                    for (int k = 0; k < 5; k++) if (k >= i && k >= j) acc += Li[k][i] * Li[k][j];
                    Sinv[i][j] = acc;
                }
            // M^-1 goes to shared memory in SOLVER-SLOT order: slot r = the r-th motor row of Bullet's row order, so the
            // sweeps below address rows, velocities and impulses with compile-time indices (registers) and only the matrix
            // through shared memory ([entry][thread]: conflict-free). A 14 x 14 local-memory array here cost 20x (r02b).
            {
                R W[5][EXO_NPRIS], U[5][EXO_NPRIS];
#pragma unroll
                for (int k = 0; k < 5; k++)
#pragma unroll
                    for (int d = 0; d < EXO_NPRIS; d++) {
                        R acc = 0;
#pragma unroll
                        for (int l = 0; l < 5; l++) acc += Sinv[k][l] * B[l][d];
                        W[k][d] = acc / D.m_d[d];
                        U[k][d] = B[k][d] / D.m_d[d];
                    }
#pragma unroll
                for (int k = 0; k < 5; k++) {
                    const int sk = D.dof_slot[k];
#pragma unroll
                    for (int l = 0; l < 5; l++) sM[D.tri[sk][D.dof_slot[l]] * NT] = Sinv[k][l];
#pragma unroll
                    for (int d = 0; d < EXO_NPRIS; d++) {
                        const int sd = D.dof_slot[5 + d];
                        sM[D.tri[sk][sd] * NT] = -W[k][d];
                        sM[D.tri[sd][sk] * NT] = -W[k][d];
                    }
                }
#pragma unroll
                for (int d = 0; d < EXO_NPRIS; d++)
#pragma unroll
                    for (int d2 = 0; d2 < EXO_NPRIS; d2++) {
                        R acc = (d == d2) ? R(1) / D.m_d[d] : R(0);
#pragma unroll
                        for (int k = 0; k < 5; k++) acc += U[k][d] * W[k][d2];
                        sM[D.tri[D.dof_slot[5 + d]][D.dof_slot[5 + d2]] * NT] = acc;
                    }
            }
            // physical DoF order -> slot order through the scratch column (static writes, uniform-address reads)
#pragma unroll
            for (int i = 0; i < EXO_NDOF; i++) sV[i * NT] = bias[i];
            R bs[EXO_NDOF], qds[EXO_NDOF];
#pragma unroll
            for (int r = 0; r < EXO_NDOF; r++) bs[r] = sV[D.slot_dof[r] * NT];
#pragma unroll
            for (int i = 0; i < EXO_NDOF; i++) sV[i * NT] = qd[i];
#pragma unroll
            for (int r = 0; r < EXO_NDOF; r++) qds[r] = sV[D.slot_dof[r] * NT];
            // ---- 5. unconstrained velocities (slot order), then one row per entry of Bullet's constraint list
            R qdu[EXO_NDOF];
#pragma unroll
            for (int r = 0; r < EXO_NDOF; r++) {
                R acc = 0;
#pragma unroll
                for (int c2 = 0; c2 < EXO_NDOF; c2++) acc -= sM[tri_c(r, c2) * NT] * bs[c2];
                qdu[r] = qds[r] + D.dt * acc;
            }
            bool has_limit = false;
#pragma unroll
            for (int j = 0; j < 5; j++) has_limit |= (q[j] - C.lower[j] <= R(0)) || (C.upper[j] - q[j] <= R(0));
            const unsigned amask0 = __activemask();
            if (!__any_sync(amask0, has_limit)) {
                // ---- 6m. motor rows only, slot order, all indices static
#pragma unroll
                for (int j = 0; j < 5; j++) sV[(5 + j) * NT] = D.kp * (tg[j] - q[j]) / D.dt;
                R dinv[EXO_NDOF], rhs[EXO_NDOF], lam[EXO_NDOF], dv[EXO_NDOF], lim[EXO_NDOF];
                const R sthr = sqrt_r(D.threshold);
#pragma unroll
                for (int r = 0; r < EXO_NDOF; r++) {
                    const int dof = D.slot_dof[r];
                    const R vu = qdu[r];
                    const R desired = dof < 5 ? sV[(5 + dof) * NT] + vu + D.kd * (R(0) - vu) : vu + R(1) * (R(0) - vu);
                    dinv[r] = R(1) / sM[tri_c(r, r) * NT];
                    rhs[r] = (desired - vu) * dinv[r];
                    lim[r] = sthr * dinv[r];
                    lam[r] = 0; dv[r] = 0;
                }
                bool going = true;
                for (int it = 0; it < D.iterations; it++) {
                    if (going) {
                        R worst = 0;
                        bool big = false;
                        if (it & 1) pgs_sweep_motors<R, true>(sM, D, dinv, rhs, lim, lam, dv, worst, big);
                        else pgs_sweep_motors<R, false>(sM, D, dinv, rhs, lim, lam, dv, worst, big);
                        sweeps_done++;
                        if (sizeof(R) == 8 ? (worst <= D.threshold) : !big) going = false;
                    }
                    if (!__any_sync(amask0, going)) break;
                }
#pragma unroll
                for (int r = 0; r < EXO_NDOF; r++) sV[r * NT] = qdu[r] + dv[r];      // new velocities, slot order
            } else
            if (D.n_cons == EXO_NCONS) {
                // per-entry data through the scratch column (static writes, warp-uniform-address reads; no register is
                // indexed at run time): first q and kp (target - q) / dt of the five arm joints, then the unconstrained velocities
#pragma unroll
                for (int j = 0; j < 5; j++) { sV[j * NT] = q[j]; sV[(5 + j) * NT] = D.kp * (tg[j] - q[j]) / D.dt; }
                R aux[EXO_NCONS];                                    // motor entry of an arm joint: its target term; limit entry: its joint angle
#pragma unroll
                for (int ci = 0; ci < EXO_NCONS; ci++) {
                    const int cc = D.cons[ci];
                    aux[ci] = cc >= 0 ? (cc < 5 ? sV[(5 + cc) * NT] : R(0)) : sV[(-cc - 1) * NT];
                }
#pragma unroll
                for (int r = 0; r < EXO_NDOF; r++) sV[r * NT] = qdu[r];
                R dinv[EXO_NCONS], rhs[EXO_NCONS], lam[EXO_NCONS], dvc[EXO_NCONS];
                unsigned act = 0, neg = 0;
#pragma unroll
                for (int ci = 0; ci < EXO_NCONS; ci++) {
                    const int cc = D.cons[ci];
                    const int dof = cc >= 0 ? cc : -cc - 1, sr = D.dof_slot[dof];
                    const R vu = sV[sr * NT];
                    dinv[ci] = R(1) / sM[D.tri[sr][sr] * NT];
                    lam[ci] = 0; dvc[ci] = 0;
                    if (cc >= 0) {
                        const R desired = cc < 5 ? aux[ci] + vu + D.kd * (R(0) - vu) : vu + R(1) * (R(0) - vu);
                        rhs[ci] = (desired - vu) * dinv[ci];
                    } else {                                 // rows only while violated; lower and upper cannot both be
                        const R pen_lo = aux[ci] - C.lower[dof], pen_hi = C.upper[dof] - aux[ci];
                        rhs[ci] = 0;
                        if (pen_lo <= R(0)) { act |= 1u << ci; rhs[ci] = (-pen_lo * D.erp / D.dt - vu) * dinv[ci]; }
                        else if (pen_hi <= R(0)) { act |= 1u << ci; neg |= 1u << ci; rhs[ci] = (-pen_hi * D.erp / D.dt + vu) * dinv[ci]; }
                    }
                }
                // ---- 6a. projected Gauss-Seidel, everything but M^-1 in registers
                const unsigned amask = __activemask();
                bool going = true;
                for (int it = 0; it < D.iterations; it++) {
                    if (going) {
                        R worst = 0;
                        // sweep direction alternates: even sweeps run the rows backwards (Bullet: iteration & 1 ? j : size-1-j);
                        // both directions are separately unrolled so the row index stays a compile-time constant
                        if (it & 1) pgs_sweep<R, true>(sM, D, act, neg, dinv, rhs, lam, dvc, worst);
                        else pgs_sweep<R, false>(sM, D, act, neg, dinv, rhs, lam, dvc, worst);
                        sweeps_done++;
                        if (worst <= D.threshold) going = false;
                    }
                    if (!__any_sync(amask, going)) break;
                }
#pragma unroll
                for (int ci = 0; ci < EXO_NCONS; ci++)               // new velocities, slot order: the scratch column still holds the
                    if (D.cons[ci] >= 0) sV[D.dof_slot[(int)D.cons[ci]] * NT] += dvc[ci];   // unconstrained ones; motor entries cover every joint once
            } else {
                // ---- 6b. a constraint list of another length (a different URDF): the general row list, indices at run time
                //      (local memory). Same arithmetic, not tuned.
#pragma unroll
                for (int i = 0; i < EXO_NDOF; i++) sV[i * NT] = i < 5 ? D.kp * (tg[i] - q[i]) / D.dt : R(0);
                int rj[EXO_MAX_ROWS];
                R rsg[EXO_MAX_ROWS], rrhs[EXO_MAX_ROWS], rdinv[EXO_MAX_ROWS], rlo[EXO_MAX_ROWS], rhi[EXO_MAX_ROWS], rlam[EXO_MAX_ROWS];
                R dvv[EXO_NDOF], qduv[EXO_NDOF];
                for (int r = 0; r < EXO_NDOF; r++) { dvv[r] = 0; qduv[r] = qdu[r]; }
                int m = 0;
                for (int ci = 0; ci < D.n_cons; ci++) {
                    const int cc = D.cons[ci];
                    if (cc >= 0) {                          // motor row of dof cc -> slot
                        const int r = D.dof_slot[cc];
                        const R vu = qduv[r], di = R(1) / sM[D.tri[r][r] * NT];
                        const R desired = cc < 5 ? sV[cc * NT] + vu + D.kd * (R(0) - vu) : vu + R(1) * (R(0) - vu);
                        rj[m] = r; rsg[m] = R(1); rdinv[m] = di; rrhs[m] = (desired - vu) * di;
                        const R cap = cc < 5 ? D.cap_pos : D.cap_vel;
                        rlo[m] = -cap; rhi[m] = cap; rlam[m] = 0; m++;
                    } else {                                // joint-limit constraint of a revolute joint: rows only while violated
                        const int j = -cc - 1, r = D.dof_slot[j];
                        const R pen_lo = q[j] - C.lower[j], pen_hi = C.upper[j] - q[j];
                        for (int sgi = 0; sgi < 2; sgi++) {
                            const R pen = sgi ? pen_hi : pen_lo, sg = sgi ? R(-1) : R(1);
                            if (pen > R(0)) continue;
                            const R di = R(1) / sM[D.tri[r][r] * NT];
                            rj[m] = r; rsg[m] = sg; rdinv[m] = di;
                            rrhs[m] = (-pen * D.erp / D.dt - sg * qduv[r]) * di;
                            rlo[m] = 0; rhi[m] = D.limit_cap; rlam[m] = 0; m++;
                        }
                    }
                }
                for (int it = 0; it < D.iterations; it++) {
                    R worst = 0;
                    for (int k = 0; k < m; k++) {
                        const int idx = (it & 1) ? k : m - 1 - k;
                        const int r = rj[idx];
                        R dl = rrhs[idx] - rsg[idx] * dvv[r] * rdinv[idx];
                        const R sum = rlam[idx] + dl;
                        if (sum < rlo[idx]) { dl = rlo[idx] - rlam[idx]; rlam[idx] = rlo[idx]; }
                        else if (sum > rhi[idx]) { dl = rhi[idx] - rlam[idx]; rlam[idx] = rhi[idx]; }
                        else rlam[idx] = sum;
                        const R imp = rsg[idx] * dl;
                        for (int i = 0; i < EXO_NDOF; i++) dvv[i] += sM[D.tri[i][r] * NT] * imp;
                        const R res = dl / rdinv[idx];
                        worst = fmax(worst, res * res);
                    }
                    sweeps_done++;
                    if (worst <= D.threshold) break;
                }
                for (int r = 0; r < EXO_NDOF; r++) sV[r * NT] = qduv[r] + dvv[r];
            }
            // ---- 7. back to the physical order, integrate, write back
            if (lag) {
#pragma unroll
                for (int d = 0; d < EXO_NPRIS; d++) dy[W_QPLAG + d] = qp[d];
            }
#pragma unroll
            for (int i = 0; i < EXO_NDOF; i++) qd[i] = sV[D.dof_slot[i] * NT];
#pragma unroll
            for (int j = 0; j < 5; j++) *exo_word(S.state, W_Q + j, e, st) = q[j] + D.dt * qd[j];
#pragma unroll
            for (int d = 0; d < EXO_NPRIS; d++) qp[d] += D.dt * qd[5 + d];
            if (!lag) {
#pragma unroll
                for (int d = 0; d < EXO_NPRIS; d++) dy[W_QPLAG + d] = qp[d];
            }
#pragma unroll
            for (int v = 0; v < NV_DYN; v++) st4(S.dyn + (size_t)v * st + e, dy + 4 * v);
        }
    }
    // solver statistics: slots 14 (sweeps) and 15 (env-steps through this kernel)
    const unsigned full = 0xffffffffu;
    const int t_sw = __reduce_add_sync(full, sweeps_done), t_n = __reduce_add_sync(full, did);
    if ((threadIdx.x & 31) == 0 && t_n) { atomicAdd(S.stats + 14, (double)t_sw); atomicAdd(S.stats + 15, (double)t_n); }
}

}  // namespace exo
