/* exo_oracle.h — CPU (fp64, plain C) restatement of the reference's ExoskeletonEnv_train hot path.
 *
 * TEST INFRASTRUCTURE ONLY. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker or the timed CPU baseline.
 * The product (libexo_env.so) never links or calls it.
 *
 * Pinning: tests/test_oracle_vs_reference.py checks every function here against
 * tests/golden/ref_*.npz, which oracle/gen_golden.py produced by running the reference's own
 * Python unmodified (numpy legacy RNG, scipy RK45). The pybullet boundary inside those goldens is
 * oracle/minibullet.py -> PARITY UNPINNED at that boundary (see its header and DESIGN.md).
 */
#ifndef EXO_ORACLE_H
#define EXO_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_NLINKS 19          /* exo_v3.urdf: 5 revolute + 14 prismatic joints */
#define ORC_MAXLEN 512         /* >= longest reference motion (347 rows) */

/* Kinematic tree in Bullet numbering (pre-order DFS, base=-1); filled by the Python side from
 * an independent URDF parse (oracle/minibullet.py). */
typedef struct {
    int    parent[ORC_NLINKS];
    int    joint_type[ORC_NLINKS];      /* 0 revolute, 1 prismatic */
    double origin_xyz[ORC_NLINKS][3];
    double origin_R[ORC_NLINKS][9];     /* row-major */
    double axis[ORC_NLINKS][3];
    double com_xyz[ORC_NLINKS][3];
    double lower[ORC_NLINKS], upper[ORC_NLINKS];
    double base_pos[3];
} OrcGeometry;

/* Constructor arguments of ExoskeletonEnv_train (Environment/Exoskeleton_env.py:38-48) + boundary knobs. */
typedef struct {
    double tremor_amplitude_range[2];
    double first_harmonics_interval[2];
    double second_harmonics_interval[2];
    double max_force_shoulder, max_force_elbow;
    double dr_actuator_end_pos_shift, dr_actuator_range, matrix_noise_fraction;
    int    tremor_sequence[7];
    int    link_state_lag;    /* minibullet knob, default 1 */
    int    joint_limits;      /* 0 off, 1 clamp */
    int    ode_mode;          /* 0 = scipy RK45 restated (reference behaviour), 1 = exact propagator */
} OrcConfig;

typedef struct OrcEnv OrcEnv;

/* np.random.seed(seed) + ExoskeletonEnv_train(...) : consumes one rand() then runs initialize_movement. */
OrcEnv* orc_create(const OrcConfig* cfg, const OrcGeometry* geo,
                   const double* motion /* [5][len]: elbow_y, elbow_z, shoulder_x, shoulder_y, shoulder_z (deg) */,
                   int len, uint32_t seed);
void orc_destroy(OrcEnv* e);
/* env.reset() -> obs[80]; returns score (=2). */
int  orc_reset(OrcEnv* e, float* obs);
/* env.step(action) ; info = 5x7 doubles in the order actuator_torques, torque_val, ampl_val,
 * tremor_torque_val, tremor_ampl_val ; rew_parts = unwanted, torque, axis, control, smoothness. */
void orc_step(OrcEnv* e, const double* action, float* obs, double* reward, int* done,
              double* info /*[35] or NULL*/, double* rew_parts /*[5] or NULL*/);
/* episode parameter views (for bit-level checks) */
void orc_get_episode(const OrcEnv* e, double* tt /*[7][len]*/, double* I, double* D, double* S,
                     double* shift /*[14][3]*/, double* fmax /*[2]: shoulder, elbow*/, double* mag);
void orc_get_bullet_q(const OrcEnv* e, double* q5);
int  orc_get_counts(const OrcEnv* e);
long orc_get_limit_crossings(const OrcEnv* e);
/* caller-side per-step metric (Exoskeleton_agent_train.py:158-170, calculate_arm_end_effector_points.py:18-50) */
void orc_forward_kinematics(const double* theta7, double L1, double L2, double L3, double* p3);
double orc_amplitude_metric(const OrcEnv* e, const double* ampl_val, const double* tremor_ampl_val,
                            double L1, double L2, double L3);
/* raw pieces exposed for unit pinning */
void orc_solve_diff_eq_rk45(const double* I, const double* D, const double* S, const double* T, double dt,
                            double* q7, int* nfev);
void orc_solve_diff_eq_exact(const double* I, const double* D, const double* S, const double* T, double dt,
                             double* q7);
void orc_propagator(const double* I, const double* D, const double* S, double dt, double* G49);
void orc_mt_seed(uint32_t* state625, uint32_t seed);
double orc_mt_random_sample(uint32_t* state625);
/* batched stepping for the timed CPU baseline: OpenMP over envs, `threads` <= 0 -> all cores */
void orc_step_many(OrcEnv** envs, int n, const double* actions /*[n][7]*/, float* obs /*[n][80]*/,
                   double* reward, int* done, int auto_reset, int threads);
int  orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
