/* exo_oracle.c — see exo_oracle.h. TEST INFRASTRUCTURE ONLY (CPU oracle, fp64, plain C).
 *
 * Every function cites the reference file:line it restates (paths relative to the reference repo).
 * Third-party arithmetic that is not under /root/reference is restated from its published algorithm:
 *   numpy legacy RandomState (MT19937; numpy 1.23.5 pinned in README.md:55-61, stream is version-stable)
 *   scipy.integrate.solve_ivp RK45 (scipy 1.11.4 pinned; scipy/integrate/_ivp/rk.py, common.py)
 *   pybullet 3.2.5 (only through the closed-form model of oracle/minibullet.py — PARITY UNPINNED)
 */
#include "exo_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

/* ------------------------------------------------------------------------------------------
 * numpy legacy RNG: np.random.seed(int) -> init_genrand; random_sample = (a>>5, b>>6) 53-bit;
 * uniform(lo,hi) = lo + (hi-lo)*random_sample; choice([-1,1],size) -> randint(0,2) -> masked
 * rejection on one 32-bit word per draw with mask 1 (never rejects).
 * ------------------------------------------------------------------------------------------ */
#define MT_N 624
#define MT_M 397

void orc_mt_seed(uint32_t* st, uint32_t seed) {
    st[0] = seed;
    for (int i = 1; i < MT_N; i++) st[i] = 1812433253u * (st[i - 1] ^ (st[i - 1] >> 30)) + (uint32_t)i;
    st[MT_N] = MT_N; /* position: force a twist on first draw */
}

static uint32_t mt_next32(uint32_t* st) {
    if (st[MT_N] >= MT_N) {
        int i;
        uint32_t y;
        for (i = 0; i < MT_N - MT_M; i++) {
            y = (st[i] & 0x80000000u) | (st[i + 1] & 0x7fffffffu);
            st[i] = st[i + MT_M] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        for (; i < MT_N - 1; i++) {
            y = (st[i] & 0x80000000u) | (st[i + 1] & 0x7fffffffu);
            st[i] = st[i + (MT_M - MT_N)] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        y = (st[MT_N - 1] & 0x80000000u) | (st[0] & 0x7fffffffu);
        st[MT_N - 1] = st[MT_M - 1] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        st[MT_N] = 0;
    }
    uint32_t y = st[st[MT_N]++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return y;
}

double orc_mt_random_sample(uint32_t* st) {
    uint32_t a = mt_next32(st) >> 5, b = mt_next32(st) >> 6;
    return (a * 67108864.0 + b) / 9007199254740992.0;
}

static double mt_uniform(uint32_t* st, double lo, double hi) { return lo + (hi - lo) * orc_mt_random_sample(st); }
static double mt_sign(uint32_t* st) { return (mt_next32(st) & 1u) ? 1.0 : -1.0; }

/* ------------------------------------------------------------------------------------------ */
struct OrcEnv {
    OrcConfig cfg;
    OrcGeometry geo;
    uint32_t mt[MT_N + 1];
    int len;                          /* max_count */
    double motion[5][ORC_MAXLEN];     /* elbow_y, elbow_z, shoulder_x, shoulder_y, shoulder_z (deg) */
    double dt;
    /* ep_state_values (Utilities/read_txt_env.py:47-70): forces + the 4 written angle series */
    double forces[7][ORC_MAXLEN];
    double tt[7][ORC_MAXLEN];         /* tremor_torque_values */
    double tremor_magnitude;
    int tremor_axis_n;
    double I0[49], D0[49], S0[49], I[49], D[49], S[49];
    double G[49]; int G_valid;
    double max_out_sh0, max_out_el0, max_out_sh, max_out_el;
    double second_prev_action[7], prev_action[7];
    double pos_vec[7][3], prev_pos_vec[7][3];
    double ref[2][3], ref_prev[2][3];
    double max_reward;
    int counts;
    /* ExoskeletonSimModel state */
    double act_pos[14][3];            /* act11,act12,act21,act22,...,act71,act72 incl. shift */
    double sh_ref[3], el_ref[3];      /* act_*_reference_positions (stale between reads) */
    double shift[14][3];
    /* minibullet */
    double q[ORC_NLINKS], qd[ORC_NLINKS], target[5];
    int have_target;
    double cache[ORC_NLINKS][3];
    long limit_crossings;
};

/* Exoskeleton_sim_pybullet.py:48-63 : link handles in the order act11,act12,act21,act22,...,act71,act72 */
static const int ACT_HANDLE[14] = {9, 5, 12, 6, 15, 8, 17, 11, 14, 7, 18, 13, 16, 10};
#define SHOULDER_HANDLE 0
#define ELBOW_HANDLE 3

/* Utilities/differential_eq_matrices.py:41-69 (seven_by_seven) */
static const double I7[49] = {
    0.269, 0, 0, 0.076, 0, 0, -0.014,
    0, 0.196, 0.083, 0, -0.002, 0.009, 0,
    0, 0.083, 0.079, 0, 0, 0.011, 0,
    0.076, 0, 0, 0.076, 0, 0, -0.012,
    0, -0.002, 0, 0, 0.002, 0, 0,
    0, 0.009, 0.011, 0, 0, 0.003, 0,
    -0.014, 0, 0, -0.012, 0, 0, 0.003};
static const double D7[49] = {
    0.756, 0.184, 0.020, 0.187, 0, 0, 0,
    0.184, 0.383, 0.267, 0, 0, 0, 0,
    0.020, 0.267, 0.524, 0, 0, 0, 0,
    0.187, 0, 0, 0.607, 0, 0, 0,
    0, 0, 0, 0, 0.021, 0.001, 0.008,
    0, 0, 0, 0, 0.001, 0.028, -0.003,
    0, 0, 0, 0, 0.008, -0.003, 0.082};
static const double S7[49] = {
    10.80, 2.626, 0.279, 2.670, 0, 0, 0,
    2.626, 5.468, 3.821, 0, 0, 0, 0,
    0.279, 3.821, 7.486, 0, 0, 0, 0,
    2.670, 0, 0, 8.670, 0, 0, 0,
    0, 0, 0, 0, 0.756, 0.018, 0.291,
    0, 0, 0, 0, 0.018, 0.992, -0.099,
    0, 0, 0, 0, 0.291, -0.099, 2.920};

/* ------------------------------------------------------------------------------------------
 * minibullet restated in C (oracle/minibullet.py: Body.forward_kinematics, World.step_simulation)
 * ------------------------------------------------------------------------------------------ */
static void mat3_mul(const double* A, const double* B, double* C) {
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) C[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
}
static void mat3_vec(const double* A, const double* v, double* o) {
    for (int i = 0; i < 3; i++) o[i] = A[i * 3] * v[0] + A[i * 3 + 1] * v[1] + A[i * 3 + 2] * v[2];
}

static void bullet_fk(const OrcGeometry* g, const double* q, double com[ORC_NLINKS][3]) {
    double R[ORC_NLINKS][9], p[ORC_NLINKS][3];
    static const double EYE[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int i = 0; i < ORC_NLINKS; i++) {
        const double* pR = g->parent[i] < 0 ? EYE : R[g->parent[i]];
        const double* pp = g->parent[i] < 0 ? g->base_pos : p[g->parent[i]];
        double Ri[9], t[3];
        mat3_mul(pR, g->origin_R[i], Ri);
        mat3_vec(pR, g->origin_xyz[i], t);
        for (int k = 0; k < 3; k++) p[i][k] = pp[k] + t[k];
        if (g->joint_type[i] == 0) {
            /* Rodrigues about the joint axis (all axes of exo_v3.urdf are local z) */
            double a[3] = {g->axis[i][0], g->axis[i][1], g->axis[i][2]};
            double n = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
            a[0] /= n; a[1] /= n; a[2] /= n;
            double c = cos(q[i]), s = sin(q[i]);
            double K[9] = {0, -a[2], a[1], a[2], 0, -a[0], -a[1], a[0], 0}, K2[9], Rq[9];
            mat3_mul(K, K, K2);
            for (int k = 0; k < 9; k++) Rq[k] = EYE[k] + s * K[k] + (1 - c) * K2[k];
            mat3_mul(Ri, Rq, R[i]);
        } else {
            memcpy(R[i], Ri, sizeof(Ri));
            double ax[3] = {g->axis[i][0] * q[i], g->axis[i][1] * q[i], g->axis[i][2] * q[i]};
            mat3_vec(Ri, ax, t);
            for (int k = 0; k < 3; k++) p[i][k] += t[k];
        }
        mat3_vec(R[i], g->com_xyz[i], t);
        for (int k = 0; k < 3; k++) com[i][k] = p[i][k] + t[k];
    }
}

static void bullet_step(OrcEnv* e) {
    const double kp = 0.1, kd = 1.0;
    if (e->cfg.link_state_lag == 1) bullet_fk(&e->geo, e->q, e->cache);
    if (e->have_target) {
        for (int j = 0; j < 5; j++) {
            double v = kp * (e->target[j] - e->q[j]) / e->dt + e->qd[j] + kd * (0.0 - e->qd[j]);
            double qn = e->q[j] + e->dt * v;
            if (qn < e->geo.lower[j] || qn > e->geo.upper[j]) {
                e->limit_crossings++;
                if (e->cfg.joint_limits == 1) {
                    qn = fmin(fmax(qn, e->geo.lower[j]), e->geo.upper[j]);
                    v = (qn - e->q[j]) / e->dt;
                }
            }
            e->qd[j] = v;
            e->q[j] = qn;
        }
    }
    if (e->cfg.link_state_lag == 0) bullet_fk(&e->geo, e->q, e->cache);
}

/* ------------------------------------------------------------------------------------------
 * 7x7 LU with partial pivoting (np.linalg.solve -> LAPACK dgesv; Utilities/calculate_joint_angles.py:12)
 * ------------------------------------------------------------------------------------------ */
typedef struct { double lu[49]; int piv[7]; } LU7;

static void lu7_factor(const double* A, LU7* f) {
    memcpy(f->lu, A, 49 * sizeof(double));
    for (int k = 0; k < 7; k++) {
        int p = k; double best = fabs(f->lu[k * 7 + k]);
        for (int i = k + 1; i < 7; i++) if (fabs(f->lu[i * 7 + k]) > best) { best = fabs(f->lu[i * 7 + k]); p = i; }
        f->piv[k] = p;
        if (p != k) for (int j = 0; j < 7; j++) { double t = f->lu[k * 7 + j]; f->lu[k * 7 + j] = f->lu[p * 7 + j]; f->lu[p * 7 + j] = t; }
        double d = f->lu[k * 7 + k];
        for (int i = k + 1; i < 7; i++) {
            double m = f->lu[i * 7 + k] / d;
            f->lu[i * 7 + k] = m;
            for (int j = k + 1; j < 7; j++) f->lu[i * 7 + j] -= m * f->lu[k * 7 + j];
        }
    }
}
static void lu7_solve(const LU7* f, const double* b, double* x) {
    double y[7];
    memcpy(y, b, sizeof(y));
    for (int k = 0; k < 7; k++) { double t = y[k]; y[k] = y[f->piv[k]]; y[f->piv[k]] = t;
        for (int i = k + 1; i < 7; i++) y[i] -= f->lu[i * 7 + k] * y[k]; }
    for (int i = 6; i >= 0; i--) { double s = y[i]; for (int j = i + 1; j < 7; j++) s -= f->lu[i * 7 + j] * x[j]; x[i] = s / f->lu[i * 7 + i]; }
}

/* dqdt of Utilities/calculate_joint_angles.py:7-16 */
typedef struct { LU7 lu; const double *D, *S, *T; int nfev; } OdeCtx;
static void dqdt(OdeCtx* c, const double* y, double* out) {
    double rhs[7];
    for (int i = 0; i < 7; i++) {
        double dq = 0, kq = 0;
        for (int j = 0; j < 7; j++) { dq += c->D[i * 7 + j] * y[7 + j]; kq += c->S[i * 7 + j] * y[j]; }
        rhs[i] = c->T[i] - dq - kq;
        out[i] = y[7 + i];
    }
    lu7_solve(&c->lu, rhs, out + 7);
    c->nfev++;
}

static double rms_norm14(const double* x) { double s = 0; for (int i = 0; i < 14; i++) s += x[i] * x[i]; return sqrt(s) / sqrt(14.0); }

/* scipy.integrate.solve_ivp(fun, [0,dt], zeros(14)) with defaults: RK45, rtol=1e-3, atol=1e-6
 * (scipy/integrate/_ivp/rk.py: rk_step, RungeKutta._step_impl, RK45 tableau; common.py: select_initial_step) */
void orc_solve_diff_eq_rk45(const double* I, const double* D, const double* S, const double* T, double dt,
                            double* q7, int* nfev) {
    static const double C[6] = {0, 1.0 / 5, 3.0 / 10, 4.0 / 5, 8.0 / 9, 1};
    static const double A[6][5] = {
        {0, 0, 0, 0, 0},
        {1.0 / 5, 0, 0, 0, 0},
        {3.0 / 40, 9.0 / 40, 0, 0, 0},
        {44.0 / 45, -56.0 / 15, 32.0 / 9, 0, 0},
        {19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561, -212.0 / 729, 0},
        {9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656}};
    static const double B[6] = {35.0 / 384, 0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84};
    static const double E[7] = {-71.0 / 57600, 0, 71.0 / 16695, -71.0 / 1920, 17253.0 / 339200, -22.0 / 525, 1.0 / 40};
    const double rtol = 1e-3, atol = 1e-6, SAFETY = 0.9, MIN_FACTOR = 0.2, MAX_FACTOR = 10.0, err_exp = -1.0 / 5.0;
    (void)C;
    OdeCtx ctx; lu7_factor(I, &ctx.lu); ctx.D = D; ctx.S = S; ctx.T = T; ctx.nfev = 0;
    double t = 0, tb = dt, y[14] = {0}, f[14], K[7][14];
    dqdt(&ctx, y, f);
    /* select_initial_step */
    double h_abs;
    {
        double scale[14], tmp[14], d0, d1, d2, h0, h1, y1[14], f1[14];
        for (int i = 0; i < 14; i++) { scale[i] = atol + fabs(y[i]) * rtol; tmp[i] = y[i] / scale[i]; }
        d0 = rms_norm14(tmp);
        for (int i = 0; i < 14; i++) tmp[i] = f[i] / scale[i];
        d1 = rms_norm14(tmp);
        h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        h0 = fmin(h0, tb - t);
        for (int i = 0; i < 14; i++) y1[i] = y[i] + h0 * f[i];
        dqdt(&ctx, y1, f1);
        for (int i = 0; i < 14; i++) tmp[i] = (f1[i] - f[i]) / scale[i];
        d2 = rms_norm14(tmp) / h0;
        h1 = (d1 <= 1e-15 && d2 <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / fmax(d1, d2), 1.0 / 5.0);
        h_abs = fmin(fmin(100 * h0, h1), tb - t);
    }
    while (t < tb) {
        double min_step = 10 * fabs(nextafter(t, INFINITY) - t);
        if (h_abs < min_step) h_abs = min_step;
        int accepted = 0, rejected = 0;
        double t_new = t, y_new[14], f_new[14];
        while (!accepted) {
            if (h_abs < min_step) { for (int i = 0; i < 7; i++) q7[i] = NAN; if (nfev) *nfev = ctx.nfev; return; }
            double h = h_abs;
            t_new = t + h;
            if (t_new - tb > 0) t_new = tb;
            h = t_new - t;
            h_abs = fabs(h);
            memcpy(K[0], f, sizeof(f));
            for (int s = 1; s < 6; s++) {
                double ys[14];
                for (int i = 0; i < 14; i++) { double dy = 0; for (int j = 0; j < s; j++) dy += K[j][i] * A[s][j]; ys[i] = y[i] + dy * h; }
                dqdt(&ctx, ys, K[s]);
            }
            for (int i = 0; i < 14; i++) { double acc = 0; for (int j = 0; j < 6; j++) acc += K[j][i] * B[j]; y_new[i] = y[i] + h * acc; }
            dqdt(&ctx, y_new, f_new);
            memcpy(K[6], f_new, sizeof(f_new));
            double errv[14];
            for (int i = 0; i < 14; i++) {
                double acc = 0; for (int j = 0; j < 7; j++) acc += K[j][i] * E[j];
                double scale = atol + fmax(fabs(y[i]), fabs(y_new[i])) * rtol;
                errv[i] = acc * h / scale;
            }
            double en = rms_norm14(errv);
            if (en < 1) {
                double factor = (en == 0) ? MAX_FACTOR : fmin(MAX_FACTOR, SAFETY * pow(en, err_exp));
                if (rejected) factor = fmin(1.0, factor);
                h_abs *= factor; accepted = 1;
            } else { h_abs *= fmax(MIN_FACTOR, SAFETY * pow(en, err_exp)); rejected = 1; }
        }
        t = t_new; memcpy(y, y_new, sizeof(y)); memcpy(f, f_new, sizeof(f));
    }
    for (int i = 0; i < 7; i++) q7[i] = y[i];
    if (nfev) *nfev = ctx.nfev;
}

/* ------------------------------------------------------------------------------------------
 * Exact solution of the same linear ODE (not in the reference; the analytic limit RK45 approximates):
 * y' = A y + B T, y(0)=0  ->  q(dt) = G T with [e^{M dt}] of M = [[A, B],[0, 0]] (21x21), G = rows 0..6
 * of the top-right block. Scaling and squaring with a degree-18 Taylor polynomial, fp64.
 * ------------------------------------------------------------------------------------------ */
#define PN 21
static void matN_mul(const double* A, const double* B, double* C) {
    for (int i = 0; i < PN; i++)
        for (int j = 0; j < PN; j++) { double s = 0; for (int k = 0; k < PN; k++) s += A[i * PN + k] * B[k * PN + j]; C[i * PN + j] = s; }
}
void orc_propagator(const double* I, const double* D, const double* S, double dt, double* G) {
    LU7 lu; lu7_factor(I, &lu);
    double M[PN * PN]; memset(M, 0, sizeof(M));
    double col[7], x[7];
    for (int j = 0; j < 7; j++) {
        M[j * PN + 7 + j] = 1.0;                                   /* qdot */
        for (int i = 0; i < 7; i++) col[i] = S[i * 7 + j]; lu7_solve(&lu, col, x);
        for (int i = 0; i < 7; i++) M[(7 + i) * PN + j] = -x[i];   /* -I^-1 S */
        for (int i = 0; i < 7; i++) col[i] = D[i * 7 + j]; lu7_solve(&lu, col, x);
        for (int i = 0; i < 7; i++) M[(7 + i) * PN + 7 + j] = -x[i];
        for (int i = 0; i < 7; i++) col[i] = (i == j); lu7_solve(&lu, col, x);
        for (int i = 0; i < 7; i++) M[(7 + i) * PN + 14 + j] = x[i]; /* I^-1 */
    }
    double nrm = 0;
    for (int i = 0; i < PN; i++) { double r = 0; for (int j = 0; j < PN; j++) r += fabs(M[i * PN + j]); if (r > nrm) nrm = r; }
    nrm *= dt;
    int s = 0; while (nrm > 0.25 && s < 60) { nrm *= 0.5; s++; }
    double h = ldexp(dt, -s);
    double X[PN * PN], P[PN * PN], Tm[PN * PN];
    for (int i = 0; i < PN * PN; i++) { M[i] *= h; X[i] = 0; P[i] = 0; }
    for (int i = 0; i < PN; i++) { X[i * PN + i] = 1; P[i * PN + i] = 1; }
    for (int k = 1; k <= 18; k++) {
        matN_mul(P, M, Tm);
        for (int i = 0; i < PN * PN; i++) { P[i] = Tm[i] / k; X[i] += P[i]; }
    }
    for (int k = 0; k < s; k++) { matN_mul(X, X, Tm); memcpy(X, Tm, sizeof(X)); }
    for (int i = 0; i < 7; i++) for (int j = 0; j < 7; j++) G[i * 7 + j] = X[i * PN + 14 + j];
}
void orc_solve_diff_eq_exact(const double* I, const double* D, const double* S, const double* T, double dt, double* q7) {
    double G[49]; orc_propagator(I, D, S, dt, G);
    for (int i = 0; i < 7; i++) { double s = 0; for (int j = 0; j < 7; j++) s += G[i * 7 + j] * T[j]; q7[i] = s; }
}

/* ------------------------------------------------------------------------------------------
 * Utilities/generate_parkinson_tremor.py:5-21,24-28,31-73
 * ------------------------------------------------------------------------------------------ */
static void generate_joint_torques_train(OrcEnv* e) {
    const int L = e->len;
    static const double JMAX[7] = {2.5, 5, 10, 5, 5, 0.5, 0.5};            /* :59 */
    double f1 = mt_uniform(e->mt, e->cfg.first_harmonics_interval[0], e->cfg.first_harmonics_interval[1]);   /* :10 */
    double f2 = mt_uniform(e->mt, e->cfg.second_harmonics_interval[0], e->cfg.second_harmonics_interval[1]); /* :11 */
    double noise[ORC_MAXLEN], w1[ORC_MAXLEN], w2[ORC_MAXLEN];
    for (int k = 0; k < L; k++) noise[k] = orc_mt_random_sample(e->mt) * 0.001;                              /* :12 */
    /* np.linspace(0, L*dt, L): step=(stop-0)/(L-1); y=arange(L)*step; y[-1]=stop   (:15) */
    double stop = L * e->dt, step = stop / (L - 1);
    for (int k = 0; k < L; k++) {
        double t = (k == L - 1) ? stop : k * step;
        w1[k] = sin(2 * M_PI * f1 * t);                                                                       /* :18 */
        w2[k] = sin(2 * M_PI * f2 * t);                                                                       /* :19 */
    }
    for (int j = 0; j < 7; j++) {
        double a1 = pow(10.0, mt_uniform(e->mt, -5, 0) / 20);                                                 /* :25 */
        double a2 = pow(10.0, mt_uniform(e->mt, -20, -10) / 20);                                              /* :26 */
        double jmax = JMAX[j] * e->tremor_magnitude;
        double g[ORC_MAXLEN], mn = INFINITY, mx = -INFINITY;
        for (int k = 0; k < L; k++) {
            g[k] = (a1 * w1[k] + a2 * w2[k] + noise[k]) * e->cfg.tremor_sequence[j];                          /* :62-63 */
            if (g[k] < mn) mn = g[k];
            if (g[k] > mx) mx = g[k];
        }
        for (int k = 0; k < L; k++) {
            double v = (-1 + 2 * (g[k] - mn) / (mx - mn)) * jmax;                                             /* :64-65 */
            if (isnan(v) || isinf(v)) v = 0;                                                                  /* :67 */
            e->tt[j][k] = v;
        }
        for (int k = 0; k < L; k++) e->tt[j][k] *= mt_sign(e->mt);                                            /* :70 */
    }
}

/* Utilities/domain_randomization_anatomical_matrices.py:4-27 */
static void add_symmetric_noise(OrcEnv* e, const double* M, double* out) {
    double n[49], f = e->cfg.matrix_noise_fraction;
    for (int i = 0; i < 49; i++) n[i] = mt_uniform(e->mt, -f, f);
    for (int i = 0; i < 7; i++)
        for (int j = 0; j < 7; j++) out[i * 7 + j] = M[i * 7 + j] + ((n[i * 7 + j] + n[j * 7 + i]) / 2) * M[i * 7 + j];
}

/* ------------------------------------------------------------------------------------------
 * ExoskeletonSimModel (Environment/Exoskeleton_sim_pybullet.py)
 * ------------------------------------------------------------------------------------------ */
static void sim_set_joint_position(OrcEnv* e, const double* pos5) { memcpy(e->target, pos5, 5 * sizeof(double)); e->have_target = 1; } /* :109-117 */
static void sim_get_actuator_positions(OrcEnv* e) {                                                     /* :119-158 */
    for (int a = 0; a < 14; a++) for (int k = 0; k < 3; k++) e->act_pos[a][k] = e->cache[ACT_HANDLE[a]][k] + e->shift[a][k];
}
static void sim_get_actuator_pos_vect(OrcEnv* e, double out[7][3]) {                                    /* :300-336 */
    for (int i = 0; i < 7; i++) {
        const double* r = (i < 2) ? e->el_ref : e->sh_ref;
        for (int k = 0; k < 3; k++) out[i][k] = e->act_pos[2 * i + 1][k] - r[k];
    }
}
static void sim_refresh_refs(OrcEnv* e) {                                                               /* :194-195, :348-349 */
    memcpy(e->sh_ref, e->cache[SHOULDER_HANDLE], sizeof(e->sh_ref));
    memcpy(e->el_ref, e->cache[ELBOW_HANDLE], sizeof(e->el_ref));
}
static void sim_get_force_components(OrcEnv* e, const double* F, double out[7][3]) {                    /* :207-298 */
    const double sh = 5;
    for (int i = 0; i < 7; i++) {
        const double* a1 = e->act_pos[2 * i];      /* act_i1 */
        const double* a2 = e->act_pos[2 * i + 1];  /* act_i2 */
        double dx = (a2[0] + sh) - (a1[0] + sh), dy = (a2[1] + sh) - (a1[1] + sh), dz = (a2[2] + sh) - (a1[2] + sh);
        double thx = atan2(dy, dx), thy = atan2(dx, dy), thz = atan2(dz, dx);
        out[i][0] = cos(thx) * F[i]; out[i][1] = cos(thy) * F[i]; out[i][2] = cos(thz) * F[i];
    }
}
static void sim_create_dummy_shift(OrcEnv* e) {                                                         /* :98-107 */
    double r = e->cfg.dr_actuator_end_pos_shift;
    for (int a = 0; a < 14; a++) for (int k = 0; k < 3; k++) e->shift[a][k] = mt_uniform(e->mt, -r, r);
}

/* ------------------------------------------------------------------------------------------
 * ExoskeletonEnv_train (Environment/Exoskeleton_env.py)
 * ------------------------------------------------------------------------------------------ */
static void update_state_vector(OrcEnv* e, float* obs) {                                                /* :487-570 */
    int c = e->counts, o = 0;
    for (int h = 2; h >= 1; h--)
        for (int i = 0; i < 7; i++) obs[o++] = (float)(e->forces[i][c - h] / (i < 2 ? e->max_out_el0 : e->max_out_sh0));
    for (int h = 2; h >= 0; h--) {
        for (int j = 0; j < 3; j++) obs[o++] = (float)(e->tt[j][c - h] / 10.0);
        obs[o++] = (float)(e->tt[3][c - h] / 5.0);
    }
    for (int i = 0; i < 7; i++) for (int k = 0; k < 3; k++) obs[o++] = (float)e->prev_pos_vec[i][k];
    for (int i = 0; i < 7; i++) for (int k = 0; k < 3; k++) obs[o++] = (float)e->pos_vec[i][k];
    for (int i = 0; i < 2; i++) for (int k = 0; k < 3; k++) obs[o++] = (float)e->ref_prev[i][k];
    for (int i = 0; i < 2; i++) for (int k = 0; k < 3; k++) obs[o++] = (float)e->ref[i][k];
}

static void joint_targets(const OrcEnv* e, int idx, const double* sh_zyx_el_y /*deg, or NULL -> reference*/, double* pos5) {
    const double d2r = M_PI / 180;
    /* order: shoulder z, y, x, elbow y, elbow z (:230-234, :427-431); motion rows: 0 el_y,1 el_z,2 sh_x,3 sh_y,4 sh_z */
    if (sh_zyx_el_y) { for (int k = 0; k < 4; k++) pos5[k] = sh_zyx_el_y[k] * d2r; }
    else { pos5[0] = e->motion[4][idx] * d2r; pos5[1] = e->motion[3][idx] * d2r; pos5[2] = e->motion[2][idx] * d2r; pos5[3] = e->motion[0][idx] * d2r; }
    pos5[4] = e->motion[1][idx] * d2r;
}

static void initialize_movement(OrcEnv* e, float* obs) {                                                /* :193-254 */
    memset(e->forces, 0, sizeof(e->forces));                                                            /* :195 */
    double lo = e->cfg.tremor_amplitude_range[0], hi = e->cfg.tremor_amplitude_range[1];
    e->tremor_magnitude = lo + (orc_mt_random_sample(e->mt) * (hi - lo));                               /* :198-199 */
    generate_joint_torques_train(e);                                                                    /* :200-205 */
    add_symmetric_noise(e, e->I0, e->I);                                                                /* :208-210 */
    add_symmetric_noise(e, e->D0, e->D);
    add_symmetric_noise(e, e->S0, e->S);
    e->G_valid = 0;
    sim_create_dummy_shift(e);                                                                          /* :213 */
    double r = e->cfg.dr_actuator_range;
    e->max_out_sh = e->max_out_sh0 * mt_uniform(e->mt, 1 - r, 1 + r);                                   /* :216 */
    e->max_out_el = e->max_out_el0 * mt_uniform(e->mt, 1 - r, 1 + r);                                   /* :217 */
    e->counts = 2;                                                                                      /* :220 */
    double pos5[5];
    joint_targets(e, e->counts - 1, NULL, pos5); sim_set_joint_position(e, pos5);                       /* :230-234 */
    sim_get_actuator_positions(e);                                                                      /* :236 */
    sim_get_actuator_pos_vect(e, e->prev_pos_vec);                                                      /* :237 */
    sim_refresh_refs(e); memcpy(e->ref_prev[0], e->sh_ref, 24); memcpy(e->ref_prev[1], e->el_ref, 24);  /* :238 */
    joint_targets(e, e->counts, NULL, pos5); sim_set_joint_position(e, pos5);                           /* :241-245 */
    sim_get_actuator_positions(e);                                                                      /* :248 */
    sim_get_actuator_pos_vect(e, e->pos_vec);                                                           /* :249 */
    sim_refresh_refs(e); memcpy(e->ref[0], e->sh_ref, 24); memcpy(e->ref[1], e->el_ref, 24);            /* :250 */
    if (obs) update_state_vector(e, obs);                                                               /* :252 */
}

OrcEnv* orc_create(const OrcConfig* cfg, const OrcGeometry* geo, const double* motion, int len, uint32_t seed) {
    if (len < 4 || len > ORC_MAXLEN) return NULL;
    OrcEnv* e = (OrcEnv*)calloc(1, sizeof(OrcEnv));
    e->cfg = *cfg; e->geo = *geo; e->len = len; e->dt = 1.0 / 40;                                       /* :59 */
    for (int r = 0; r < 5; r++) memcpy(e->motion[r], motion + (size_t)r * len, len * sizeof(double));
    orc_mt_seed(e->mt, seed);                                                                           /* Utilities/seed_setting_.py:9 */
    (void)orc_mt_random_sample(e->mt);                                                                  /* :70 */
    e->tremor_axis_n = 0; for (int j = 0; j < 7; j++) e->tremor_axis_n += cfg->tremor_sequence[j];      /* :71 */
    e->max_out_sh0 = e->max_out_sh = cfg->max_force_shoulder; e->max_out_el0 = e->max_out_el = cfg->max_force_elbow;
    memcpy(e->I0, I7, sizeof(I7)); memcpy(e->D0, D7, sizeof(D7)); memcpy(e->S0, S7, sizeof(S7));        /* :137 */
    e->max_reward = e->tremor_axis_n * 0.5 + 0.9 + 0.05 + 0.05 + 0.5;                                   /* :161-169 */
    bullet_fk(&e->geo, e->q, e->cache);                                                                 /* loadURDF */
    initialize_movement(e, NULL);                                                                       /* :172 */
    return e;
}
void orc_destroy(OrcEnv* e) { free(e); }
int orc_reset(OrcEnv* e, float* obs) { initialize_movement(e, obs); return e->counts; }                 /* :473-478 */

static void cross3(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}

void orc_step(OrcEnv* e, const double* action, float* obs, double* reward, int* done, double* info, double* rew_parts) {
    const double eps = 1e-10;                                                                           /* :175 */
    const int c = e->counts;
    double F[7];
    for (int i = 0; i < 7; i++) F[i] = ((action[i] + 1) / 2) * (i < 2 ? e->max_out_el : e->max_out_sh); /* :256-266 */
    for (int i = 0; i < 7; i++) e->forces[i][c] = F[i];                                                 /* :373-379 */
    sim_get_actuator_positions(e);                                                                      /* :382 */
    memcpy(e->ref_prev, e->ref, sizeof(e->ref));                                                        /* :383 */
    double fc[7][3]; sim_get_force_components(e, F, fc);                                                /* :384 */
    memcpy(e->prev_pos_vec, e->pos_vec, sizeof(e->pos_vec));                                            /* :385 */
    sim_get_actuator_pos_vect(e, e->pos_vec);                                                           /* :386 (stale refs) */
    sim_refresh_refs(e); memcpy(e->ref[0], e->sh_ref, 24); memcpy(e->ref[1], e->el_ref, 24);            /* :387 */
    double tq[7][3];                                                                                    /* :388 -> :177-187, sim:192-205 */
    for (int i = 0; i < 7; i++) {
        const double* rr = (i < 2) ? e->el_ref : e->sh_ref;
        double rad[3] = {rr[0] - e->act_pos[2 * i + 1][0], rr[1] - e->act_pos[2 * i + 1][1], rr[2] - e->act_pos[2 * i + 1][2]};
        cross3(fc[i], rad, tq[i]);
    }
    double at[7] = {0};                                                                                 /* :394-400 (order 3,4,5,7,6) */
    for (int k = 0; k < 3; k++) {
        static const int comp[3] = {1, 0, 2};
        at[k] = tq[2][comp[k]] + tq[3][comp[k]] + tq[4][comp[k]] + tq[6][comp[k]] + tq[5][comp[k]];
    }
    at[3] = fabs(tq[0][1]) - fabs(tq[1][1]);
    double ttq[7], tv[7];
    for (int j = 0; j < 7; j++) { ttq[j] = e->tt[j][c]; tv[j] = ttq[j] + at[j]; }                       /* :403-406 */
    double qa[7], qt[7];
    if (e->cfg.ode_mode == 0) {
        orc_solve_diff_eq_rk45(e->I, e->D, e->S, tv, e->dt, qa, NULL);                                  /* :409 */
        orc_solve_diff_eq_rk45(e->I, e->D, e->S, ttq, e->dt, qt, NULL);                                 /* :413 */
    } else {
        if (!e->G_valid) { orc_propagator(e->I, e->D, e->S, e->dt, e->G); e->G_valid = 1; }
        for (int i = 0; i < 7; i++) { double s1 = 0, s2 = 0; for (int j = 0; j < 7; j++) { s1 += e->G[i * 7 + j] * tv[j]; s2 += e->G[i * 7 + j] * ttq[j]; } qa[i] = s1; qt[i] = s2; }
    }
    for (int i = 0; i < 7; i++) { qa[i] = qa[i] * (180 / M_PI); qt[i] = qt[i] * (180 / M_PI); }         /* :417-418 */
    double ang[4] = {e->motion[4][c] + qa[2], e->motion[3][c] + qa[0], e->motion[2][c] + qa[1], e->motion[0][c] + qa[3]}; /* :421-424 */
    double pos5[5]; joint_targets(e, c, ang, pos5); sim_set_joint_position(e, pos5);                    /* :427-431 */
    bullet_step(e);                                                                                     /* :433 */
    /* rewards :341-366 */
    double m = e->max_out_el + e->max_out_sh, n = e->tremor_axis_n;
    double sum_unw = 0, sum_tor = 0, sumF = 0, smooth = 0; int axis_cnt = 0;
    for (int j = 0; j < 4; j++) {
        if (e->cfg.tremor_sequence[j] == 0) sum_unw += fabs(tv[j]);                                     /* :331-334 */
        if (e->cfg.tremor_sequence[j] == 1) sum_tor += (fabs(tv[j]) - fabs(ttq[j])) / fabs(ttq[j]) + 1; /* :315-318 */
    }
    double r_unw = exp(-(sum_unw / (m / 4 / n)) + eps) * 0.5;                                           /* :336-338 */
    double r_tor = exp((-sum_tor + eps) / n) * 0.9;                                                     /* :320-321 */
    for (int j = 0; j < 7; j++) {                                                                       /* :301-304 */
        double red = (fabs(tv[j]) - fabs(ttq[j])) / (fabs(ttq[j]) + eps) * 100;
        if (isnan(red) || isinf(red)) red = 0;
        if (red < 0) axis_cnt++;
    }
    double r_axis = axis_cnt * 0.5;
    for (int i = 0; i < 7; i++) { sumF += F[i]; double d = F[i] - 2 * e->prev_action[i] + e->second_prev_action[i]; smooth += d * d; }
    double r_ctl = exp(-(sumF / (m / 2)) + eps) * 0.05;                                                 /* :275-277 */
    double r_smo = 0.05 * exp(-((smooth / 7) / (m / 4)) + eps);                                         /* :288-290 */
    *reward = (r_axis + r_tor + r_smo + r_ctl + r_unw) / e->max_reward;                                 /* :356 */
    e->counts += 1;                                                                                     /* :448 */
    update_state_vector(e, obs);                                                                        /* :451 */
    memcpy(e->second_prev_action, e->prev_action, sizeof(F)); memcpy(e->prev_action, F, sizeof(F));     /* :454-455 */
    *done = e->counts >= e->len - 1;                                                                    /* :457 */
    if (info) { memcpy(info, at, 56); memcpy(info + 7, tv, 56); memcpy(info + 14, qa, 56); memcpy(info + 21, ttq, 56); memcpy(info + 28, qt, 56); }
    if (rew_parts) { rew_parts[0] = r_unw; rew_parts[1] = r_tor; rew_parts[2] = r_axis; rew_parts[3] = r_ctl; rew_parts[4] = r_smo; }
}

void orc_get_episode(const OrcEnv* e, double* tt, double* I, double* D, double* S, double* shift, double* fmax, double* mag) {
    if (tt) for (int j = 0; j < 7; j++) memcpy(tt + (size_t)j * e->len, e->tt[j], e->len * sizeof(double));
    if (I) memcpy(I, e->I, sizeof(e->I));
    if (D) memcpy(D, e->D, sizeof(e->D));
    if (S) memcpy(S, e->S, sizeof(e->S));
    if (shift) memcpy(shift, e->shift, sizeof(e->shift));
    if (fmax) { fmax[0] = e->max_out_sh; fmax[1] = e->max_out_el; }
    if (mag) *mag = e->tremor_magnitude;
}
void orc_get_bullet_q(const OrcEnv* e, double* q5) { memcpy(q5, e->q, 5 * sizeof(double)); }
int orc_get_counts(const OrcEnv* e) { return e->counts; }
long orc_get_limit_crossings(const OrcEnv* e) { return e->limit_crossings; }

/* Utilities/calculate_arm_end_effector_points.py:8-15,18-50 : D-H chain A1..A7, end effector = T[:3,3] */
void orc_forward_kinematics(const double* th, double L1, double L2, double L3, double* p3) {
    static const double alpha[7] = {M_PI / 2, M_PI / 2, -M_PI / 2, M_PI / 2, M_PI / 2, M_PI / 2, M_PI / 2};
    const double d[7] = {0, 0, L1, 0, L2, 0, 0};
    (void)L3;
    double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, A[16], R[16];
    for (int k = 0; k < 7; k++) {
        double ct = cos(th[k]), st = sin(th[k]), ca = cos(alpha[k]), sa = sin(alpha[k]);
        double a = 0;
        double Ak[16] = {ct, -st * ca, st * sa, a * ct, st, ct * ca, -ct * sa, a * st, 0, sa, ca, d[k], 0, 0, 0, 1};
        memcpy(A, Ak, sizeof(A));
        for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { double s = 0; for (int l = 0; l < 4; l++) s += T[i * 4 + l] * A[l * 4 + j]; R[i * 4 + j] = s; }
        memcpy(T, R, sizeof(T));
    }
    p3[0] = T[3]; p3[1] = T[7]; p3[2] = T[11];
}

/* Exoskeleton_agent_train.py:158-170 with return_original_joint_angles (Exoskeleton_env.py:580-592) at the
 * env's CURRENT counts (i.e. after the step's increment, as the training script calls it). */
double orc_amplitude_metric(const OrcEnv* e, const double* ampl, const double* tampl, double L1, double L2, double L3) {
    const double d2r = M_PI / 180; int c = e->counts;
    double base[7] = {e->motion[2][c] * d2r, e->motion[3][c] * d2r, e->motion[4][c] * d2r, e->motion[0][c] * d2r, e->motion[1][c] * d2r, 0, 0};
    double s[7], u[7], p0[3], p1[3], p2[3];
    for (int i = 0; i < 7; i++) { s[i] = ampl[i] * d2r + base[i]; u[i] = tampl[i] * d2r + base[i]; }
    orc_forward_kinematics(base, L1, L2, L3, p0); orc_forward_kinematics(s, L1, L2, L3, p1); orc_forward_kinematics(u, L1, L2, L3, p2);
    double ds = 0, du = 0;
    for (int k = 0; k < 3; k++) { ds += (p1[k] - p0[k]) * (p1[k] - p0[k]); du += (p2[k] - p0[k]) * (p2[k] - p0[k]); }
    ds = sqrt(ds); du = sqrt(du);
    return ((ds - du) / du) * 100;
}

int orc_max_threads(void) { long n = sysconf(_SC_NPROCESSORS_ONLN); return n > 0 ? (int)n : 1; }

typedef struct { OrcEnv** envs; int lo, hi; const double* actions; float* obs; double* reward; int* done; int auto_reset; } ManyJob;
static void* many_worker(void* p) {
    ManyJob* j = (ManyJob*)p;
    for (int i = j->lo; i < j->hi; i++) {
        orc_step(j->envs[i], j->actions + (size_t)i * 7, j->obs + (size_t)i * 80, j->reward + i, j->done + i, NULL, NULL);
        if (j->auto_reset && j->done[i]) orc_reset(j->envs[i], j->obs + (size_t)i * 80);
    }
    return NULL;
}
/* one pass over n independent envs, split in contiguous slices over `threads` pthreads */
void orc_step_many(OrcEnv** envs, int n, const double* actions, float* obs, double* reward, int* done, int auto_reset, int threads) {
    if (threads <= 0) threads = orc_max_threads();
    if (threads > n) threads = n > 0 ? n : 1;
    if (threads > 256) threads = 256;
    pthread_t tid[256]; ManyJob job[256];
    for (int t = 0; t < threads; t++) {
        ManyJob jb = {envs, (int)((long)n * t / threads), (int)((long)n * (t + 1) / threads), actions, obs, reward, done, auto_reset};
        job[t] = jb;
        if (t > 0) pthread_create(&tid[t], NULL, many_worker, &job[t]);
    }
    many_worker(&job[0]);
    for (int t = 1; t < threads; t++) pthread_join(tid[t], NULL);
}
