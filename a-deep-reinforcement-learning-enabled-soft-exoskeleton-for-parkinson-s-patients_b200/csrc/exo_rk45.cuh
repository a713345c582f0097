// exo_rk45.cuh — the reference's OWN integrator for the arm ODE, fp64, for the verification mode ExoConfig.ode = EXO_ODE_RK45.
//
// Utilities/calculate_joint_angles.py:5-22 integrates  y' = [q'; I^-1 (T - D q' - S q)],  y in R^14, from y(0) = 0 over
// [0, dt] with scipy.integrate.solve_ivp at its defaults: explicit Runge-Kutta 5(4) (Dormand-Prince), rtol 1e-3, atol 1e-6,
// scipy's initial-step heuristic and step controller (scipy/integrate/_ivp/rk.py: rk_step, RungeKutta._step_impl;
// common.py: select_initial_step; constants SAFETY 0.9, MIN_FACTOR 0.2, MAX_FACTOR 10, error exponent -1/5, RMS norm).
// np.linalg.solve(I, .) inside the right-hand side is an LU factorisation with partial pivoting (LAPACK dgesv), done once
// here. The product path replaces all of this by the exact solution q(dt) = G T (exo_reset.cuh); this file exists so that
// the CUDA path can be compared with the reference's goldens WITHOUT the 1e-3 truncation error RK45 itself carries:
// tests/test_gpu_parity.py holds ampl_val to 1e-10 in this mode. One thread per env, local memory, ~40 right-hand-side
// evaluations per solve, two solves per step: a debug instantiation (fp64 handles only), not a throughput path.
#pragma once
#include <math.h>

namespace exo {

struct Lu7 { double lu[49]; int piv[7]; };

__device__ inline void lu7_factor(const double* A, Lu7& f) {
    for (int i = 0; i < 49; i++) f.lu[i] = A[i];
    for (int k = 0; k < 7; k++) {
        int p = k; double best = fabs(f.lu[k * 7 + k]);
        for (int i = k + 1; i < 7; i++) if (fabs(f.lu[i * 7 + k]) > best) { best = fabs(f.lu[i * 7 + k]); p = i; }
        f.piv[k] = p;
        if (p != k) for (int j = 0; j < 7; j++) { const double t = f.lu[k * 7 + j]; f.lu[k * 7 + j] = f.lu[p * 7 + j]; f.lu[p * 7 + j] = t; }
        const double d = f.lu[k * 7 + k];
        for (int i = k + 1; i < 7; i++) {
            const double m = f.lu[i * 7 + k] / d;
            f.lu[i * 7 + k] = m;
            for (int j = k + 1; j < 7; j++) f.lu[i * 7 + j] -= m * f.lu[k * 7 + j];
        }
    }
}
__device__ inline void lu7_solve(const Lu7& f, const double* b, double* x) {
    double y[7];
    for (int i = 0; i < 7; i++) y[i] = b[i];
    for (int k = 0; k < 7; k++) {
        const double t = y[k]; y[k] = y[f.piv[k]]; y[f.piv[k]] = t;
        for (int i = k + 1; i < 7; i++) y[i] -= f.lu[i * 7 + k] * y[k];
    }
    for (int i = 6; i >= 0; i--) { double s = y[i]; for (int j = i + 1; j < 7; j++) s -= f.lu[i * 7 + j] * x[j]; x[i] = s / f.lu[i * 7 + i]; }
}
// dqdt of calculate_joint_angles.py:7-16
__device__ inline void rk45_rhs(const Lu7& lu, const double* D, const double* S, const double* T, const double* y, double* out) {
    double rhs[7];
    for (int i = 0; i < 7; i++) {
        double dq = 0, kq = 0;
        for (int j = 0; j < 7; j++) { dq += D[i * 7 + j] * y[7 + j]; kq += S[i * 7 + j] * y[j]; }
        rhs[i] = T[i] - dq - kq;
        out[i] = y[7 + i];
    }
    lu7_solve(lu, rhs, out + 7);
}
__device__ inline double rms14(const double* x) { double s = 0; for (int i = 0; i < 14; i++) s += x[i] * x[i]; return sqrt(s) / sqrt(14.0); }

// solve_ivp(dqdt, [0, dt], zeros(14)).y[:7, -1]; ids = I | D | S (3 x 49 doubles)
static __device__ __noinline__ void rk45_solve(const double* ids, const double* T, double dt, double* q7) {
    const double A[6][5] = {
        {0, 0, 0, 0, 0},
        {1.0 / 5, 0, 0, 0, 0},
        {3.0 / 40, 9.0 / 40, 0, 0, 0},
        {44.0 / 45, -56.0 / 15, 32.0 / 9, 0, 0},
        {19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561, -212.0 / 729, 0},
        {9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656}};
    const double B[6] = {35.0 / 384, 0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84};
    const double E[7] = {-71.0 / 57600, 0, 71.0 / 16695, -71.0 / 1920, 17253.0 / 339200, -22.0 / 525, 1.0 / 40};
    const double rtol = 1e-3, atol = 1e-6, SAFETY = 0.9, MIN_FACTOR = 0.2, MAX_FACTOR = 10.0, err_exp = -1.0 / 5.0;
    const double* D = ids + 49;
    const double* S = ids + 98;
    Lu7 lu;
    lu7_factor(ids, lu);
    double t = 0, tb = dt, y[14], f[14], K[7][14];
    for (int i = 0; i < 14; i++) y[i] = 0;
    rk45_rhs(lu, D, S, T, y, f);
    double h_abs;
    {   // select_initial_step
        double scale[14], tmp[14], y1[14], f1[14];
        for (int i = 0; i < 14; i++) { scale[i] = atol + fabs(y[i]) * rtol; tmp[i] = y[i] / scale[i]; }
        const double d0 = rms14(tmp);
        for (int i = 0; i < 14; i++) tmp[i] = f[i] / scale[i];
        const double d1 = rms14(tmp);
        double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        h0 = fmin(h0, tb - t);
        for (int i = 0; i < 14; i++) y1[i] = y[i] + h0 * f[i];
        rk45_rhs(lu, D, S, T, y1, f1);
        for (int i = 0; i < 14; i++) tmp[i] = (f1[i] - f[i]) / scale[i];
        const double d2 = rms14(tmp) / h0;
        const double h1 = (d1 <= 1e-15 && d2 <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / fmax(d1, d2), 1.0 / 5.0);
        h_abs = fmin(fmin(100 * h0, h1), tb - t);
    }
    while (t < tb) {
        const double min_step = 10 * fabs(__longlong_as_double(__double_as_longlong(t) + 1) - t);   // 10 |nextafter(t, inf) - t|, t >= 0
        if (h_abs < min_step) h_abs = min_step;
        bool accepted = false, rejected = false;
        double t_new = t, y_new[14], f_new[14];
        while (!accepted) {
            if (h_abs < min_step) { for (int i = 0; i < 7; i++) q7[i] = NAN; return; }
            double h = h_abs;
            t_new = t + h;
            if (t_new - tb > 0) t_new = tb;
            h = t_new - t;
            h_abs = fabs(h);
            for (int i = 0; i < 14; i++) K[0][i] = f[i];
            for (int s = 1; s < 6; s++) {
                double ys[14];
                for (int i = 0; i < 14; i++) { double dy = 0; for (int j = 0; j < s; j++) dy += K[j][i] * A[s][j]; ys[i] = y[i] + dy * h; }
                rk45_rhs(lu, D, S, T, ys, K[s]);
            }
            for (int i = 0; i < 14; i++) { double acc = 0; for (int j = 0; j < 6; j++) acc += K[j][i] * B[j]; y_new[i] = y[i] + h * acc; }
            rk45_rhs(lu, D, S, T, y_new, f_new);
            for (int i = 0; i < 14; i++) K[6][i] = f_new[i];
            double errv[14];
            for (int i = 0; i < 14; i++) {
                double acc = 0; for (int j = 0; j < 7; j++) acc += K[j][i] * E[j];
                const double scale = atol + fmax(fabs(y[i]), fabs(y_new[i])) * rtol;
                errv[i] = acc * h / scale;
            }
            const double en = rms14(errv);
            if (en < 1) {
                double factor = (en == 0) ? MAX_FACTOR : fmin(MAX_FACTOR, SAFETY * pow(en, err_exp));
                if (rejected) factor = fmin(1.0, factor);
                h_abs *= factor; accepted = true;
            } else { h_abs *= fmax(MIN_FACTOR, SAFETY * pow(en, err_exp)); rejected = true; }
        }
        t = t_new;
        for (int i = 0; i < 14; i++) { y[i] = y_new[i]; f[i] = f_new[i]; }
    }
    for (int i = 0; i < 7; i++) q7[i] = y[i];
}

}  // namespace exo
