/* exo_env.h — C-ABI of libexo_env.so: the batched, B200-resident replacement of the reference's
 * `ExoskeletonEnv_train` reset()/step() hot path (Environment/Exoskeleton_env.py:34-605 and everything it
 * calls in Environment/Exoskeleton_sim_pybullet.py and Utilities/*.py; map in DESIGN.md).
 *
 * Conventions
 *  - plain C types only; one handle per GPU; a handle is not thread-safe.
 *  - every function returns 0 on success or a negative ExoStatus; exo_last_error() gives the text
 *    (thread-local). Nothing falls back to the CPU: without a usable sm_100 device exo_create fails.
 *  - "dev" pointers are device pointers owned by the caller (e.g. torch tensors); "host" pointers are
 *    host memory. Kernels are launched on the caller-supplied cudaStream_t (passed as void*); device
 *    entry points never synchronise, host-buffer entry points synchronise that stream before returning.
 *  - `real` below is float when ExoConfig.precision == EXO_F32 and double when EXO_F64.
 *  - env arrays are row-major [n_envs, k] at this boundary; inside the handle state is structure-of-arrays.
 *  - episode parameters (tremor table, I/D/S, propagator, shifts, force maxima) are double-buffered per env and generated
 *    one episode ahead: exo_seed fills the spare slot of every env; starting an episode (exo_reset, or - with auto_reset -
 *    inside exo_step's kernel when an episode ends) flips the env to its spare slot and queues the env on a device-side
 *    work list, from which a generation pass on an internal side stream refills the vacated slot while later steps run.
 *    The draws each env makes from its private numpy-legacy MT19937 stream, and their order, are those of the reference
 *    (reset after reset); only WHEN they are computed moves. No host synchronisation anywhere.
 */
#ifndef EXO_ENV_H
#define EXO_ENV_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EXO_OBS_DIM 80      /* Exoskeleton_env.py:99-111,487-570 */
#define EXO_ACT_DIM 7       /* Exoskeleton_env.py:74 */
#define EXO_NLINKS 19       /* Simulation/exo_v3.urdf: 5 revolute + 14 prismatic joints */
#define EXO_N_STATS 16

typedef enum {
    EXO_OK = 0, EXO_ERR_ARG = -1, EXO_ERR_CUDA = -2, EXO_ERR_STATE = -3, EXO_ERR_TOPOLOGY = -4, EXO_ERR_NOMEM = -5
} ExoStatus;

enum { EXO_F32 = 0, EXO_F64 = 1 };
/* How the arm ODE of Utilities/calculate_joint_angles.py:5-22 is integrated over one dt from rest:
 *   EXO_ODE_PROPAGATOR  exact solution q(dt) = G T, G built per episode in fp64 (default; the product path);
 *   EXO_ODE_RK45        the reference's own integrator, scipy's RK45 at its default tolerances restated on the device
 *                       (csrc/exo_rk45.cuh), fp64 handles only: a verification mode that removes RK45's own 1e-3 truncation
 *                       error from comparisons with reference goldens. ~100x slower than the default. */
enum { EXO_ODE_PROPAGATOR = 0, EXO_ODE_RK45 = 1 };
/* What stands in for pybullet's stepSimulation (Exoskeleton_env.py:433) behind the POSITION_CONTROL motors
 * (Exoskeleton_sim_pybullet.py:109-117):
 *   EXO_MOTOR_IDEAL  the fixed point of Bullet's constraint solve, q+ = q + kp (target - q), inside the step kernel (default);
 *   EXO_MOTOR_FULL   the general multibody step: forward dynamics of the URDF tree (gravity, link damping, Coriolis),
 *                    motor + joint-limit rows, numSolverIterations sweeps of projected Gauss-Seidel in Bullet's row order
 *                    with its early exit, semi-implicit Euler (csrc/exo_motor.cuh, oracle/multibody.py). Needs
 *                    exo_set_motor_model before exo_seed. */
enum { EXO_MOTOR_IDEAL = 0, EXO_MOTOR_FULL = 1 };

/* Constructor arguments of ExoskeletonEnv_train (Exoskeleton_env.py:38-48), shared by all envs of a handle,
 * plus the knobs of the restated pybullet boundary (oracle/minibullet.py) and kernel options. */
typedef struct {
    double tremor_amplitude_range[2];
    double first_harmonics_interval[2];
    double second_harmonics_interval[2];
    double max_force_shoulder, max_force_elbow;
    double dr_actuator_end_pos_shift, dr_actuator_range, matrix_noise_fraction;
    double humerus_length, forearm_length, hand_length;   /* Exoskeleton_agent_train.py:37-41, amplitude metric */
    int32_t link_state_lag;   /* 1: getLinkState lags the joint state by one stepSimulation (default), 0: no lag */
    int32_t joint_limits;     /* 0: motor row wins (default), 1: clamp to the URDF limits */
    int32_t precision;        /* EXO_F32 | EXO_F64 */
    int32_t ode;              /* EXO_ODE_PROPAGATOR (default) | EXO_ODE_RK45 (fp64 verification mode) */
    int32_t auto_reset;       /* 1: exo_step re-initialises finished envs in the same call (obs = first obs of next episode) */
    int32_t reference_ctor;   /* 1: exo_seed also performs the constructor's draws: one rand() (Exoskeleton_env.py:70)
                                 and one initialize_movement (:172), so streams line up with the reference */
    int32_t metrics;          /* 1: accumulate the training script's per-step suppression metrics on device */
    int32_t motor_model;      /* EXO_MOTOR_IDEAL (default) | EXO_MOTOR_FULL */
} ExoConfig;

/* Kinematic tree of the URDF in Bullet numbering (pre-order DFS over child joints in file order, base = -1). */
typedef struct {
    int32_t parent[EXO_NLINKS];
    int32_t joint_type[EXO_NLINKS];       /* 0 revolute, 1 prismatic */
    double  origin_xyz[EXO_NLINKS][3];
    double  origin_R[EXO_NLINKS][9];      /* row-major rotation of the joint origin rpy (Rz*Ry*Rx) */
    double  axis[EXO_NLINKS][3];
    double  com_xyz[EXO_NLINKS][3];       /* inertial origin: getLinkState(...)[0] is the COM */
    double  lower[EXO_NLINKS], upper[EXO_NLINKS];
    double  base_pos[3];                  /* world position of the base link frame (loadURDF basePosition) */
    int32_t act_link[14];                 /* link handles act11,act12,act21,...,act72 (Exoskeleton_sim_pybullet.py:48-61) */
    int32_t shoulder_link, elbow_link;    /* Exoskeleton_sim_pybullet.py:62-63 */
} ExoGeometry;

/* Inertial data of the URDF links, Bullet numbering (Simulation/exo_v3.urdf <inertial> / <collision> elements). pybullet
 * recomputes link inertias from the collision shape unless URDF_USE_INERTIA_FROM_FILE is passed (the reference passes no
 * flags, Exoskeleton_sim_pybullet.py:18), so both descriptions travel. */
typedef struct {
    double mass[EXO_NLINKS];
    double inertia_diag[EXO_NLINKS][3];   /* ixx, iyy, izz of <inertia> (products of inertia are zero in this URDF) */
    double inertial_R[EXO_NLINKS][9];     /* rotation of the inertial frame in the link frame (<inertial><origin rpy>) */
    double box_size[EXO_NLINKS][3];       /* the link's <collision> box */
    double box_xyz[EXO_NLINKS][3];
    double box_R[EXO_NLINKS][9];
} ExoLinkDynamics;

/* Settings of the general multibody step; the values in comments are pybullet 3.2.5's defaults as the reference uses it. */
typedef struct {
    int32_t inertia_from_file;     /* 0: inertia of the collision box's AABB in the inertial frame (pybullet default), 1: <inertia> */
    int32_t iterations;            /* numSolverIterations: 50 */
    int32_t row_order;             /* 0: order left by Bullet's quickSort of the world's constraints by island id, 1: world order */
    int32_t reserved;
    double  residual_threshold;    /* leastSquaresResidualThreshold on the largest squared row velocity error of a sweep: 1e-7 */
    double  kp, kd, max_force;     /* setJointMotorControlArray defaults: 0.1, 1.0, 1e5 N (impulse cap = max_force * dt) */
    double  default_motor_impulse; /* the velocity motors createJointMotors puts on every joint: 1.0 N s */
    double  linear_damping, angular_damping;   /* btMultiBody per-link damping: 0.04, 0.04 */
    double  gravity[3];            /* Exoskeleton_env.py:116: 0, 0, -9.81 */
    double  erp, limit_max_impulse;            /* joint-limit rows: 0.2, 100 */
} ExoMotorModel;

/* Optional per-step outputs = the reference's `info` dict (Exoskeleton_env.py:464-469). Any pointer may be NULL. */
typedef struct {
    void* actuator_torques;   /* real [n,7] */
    void* torque_val;         /* real [n,7] */
    void* ampl_val;           /* real [n,7] degrees */
    void* tremor_torque_val;  /* real [n,7] */
    void* tremor_ampl_val;    /* real [n,7] degrees */
    void* reward_parts;       /* real [n,5]: unwanted, torque, axis, control, smoothness (Exoskeleton_env.py:358-364) */
    void* amp_metric;         /* real [n]: end-effector amplitude change % (Exoskeleton_agent_train.py:158-170) */
} ExoInfo;

/* Resident per-env state, addressed by field. The layout inside the handle (vector-major structure-of-arrays,
 * DESIGN.md) is private; exo_get_field / exo_set_field gather/scatter env-major [n_envs, k] device buffers.
 * For parity tests and checkpoint/restore. */
typedef enum {
    EXO_FIELD_Q = 0,          /* real [5]   Bullet joint positions: shoulder z,y,x, elbow y,z (rad) */
    EXO_FIELD_Q_LAG,          /* real [5]   joint positions the link-state cache was computed from */
    EXO_FIELD_POS_VEC,        /* real [21]  position_vectors (Exoskeleton_env.py:386) */
    EXO_FIELD_ELBOW_REF,      /* real [3]   act_elbow_reference_positions as last read */
    EXO_FIELD_PREV_ACTION,    /* real [7]   Newtons (Exoskeleton_env.py:455) */
    EXO_FIELD_PREV_ACTION2,   /* real [7]   (Exoskeleton_env.py:454) */
    EXO_FIELD_FORCE_MAX,      /* real [2]   domain-randomised max_output_shoulder, max_output_elbow (:216-217) */
    EXO_FIELD_SHIFT,          /* real [42]  dummy_shift_coordinates (Exoskeleton_sim_pybullet.py:103-107) */
    EXO_FIELD_G,              /* real [49]  propagator of I q'' + D q' + S q = T over one dt from rest: q(dt) = G T (rad/Nm) */
    EXO_FIELD_IDS,            /* f64  [147] I_current, D_current, S_current (Exoskeleton_env.py:208-210) */
    EXO_FIELD_TREMOR_TABLE,   /* real [table_len * table_axes] tremor_torque_values, time-major */
    EXO_FIELD_TREMOR_MAG,     /* f64  [1]   tremor_magnitude (:199) */
    EXO_FIELD_COUNTS,         /* i32  [1]   self.counts */
    EXO_FIELD_META,           /* i32  [1]   bits 0-6 tremor_sequence, bits 8-15 motion id, bits 16-31 max_count */
    EXO_FIELD_MT_STATE,       /* u32  [625] numpy legacy MT19937 key + position */
    EXO_FIELD_EP_RETURN,      /* real [1]   running episode return */
    EXO_FIELD_QD,             /* real [14]  EXO_MOTOR_FULL: joint velocities: 5 arm joints, 7 humerus dummies (act11, act21, act32,
                                            act42, act52, act62, act72), 2 forearm dummies (act12, act22) */
    EXO_FIELD_Q_PRIS,         /* real [9]   EXO_MOTOR_FULL: prismatic positions of those 9 dummies (m) */
    EXO_FIELD_Q_PRIS_LAG,     /* real [9]   EXO_MOTOR_FULL: the prismatic positions behind the link-state cache */
    EXO_N_FIELDS
} ExoField;
enum { EXO_DT_REAL = 0, EXO_DT_F64 = 1, EXO_DT_I32 = 2, EXO_DT_U32 = 3 };
/* Episode parameters are generated ONE EPISODE AHEAD (see the conventions above). OR this flag into a field id to address
 * the prefetched copy - the parameters the env's NEXT episode will use - of FORCE_MAX, SHIFT, G, IDS, TREMOR_TABLE and
 * TREMOR_MAG. EXO_FIELD_MT_STATE is the stream position AFTER the prefetched episode's draws. A checkpoint is every field
 * plus the EXO_FIELD_NEXT copies. */
#define EXO_FIELD_NEXT 0x100

typedef struct ExoHandle ExoHandle;

const char* exo_last_error(void);
int exo_abi_version(void);

/* ExoskeletonEnv_train.__init__ for n_envs environments at once (Exoskeleton_env.py:38-175, URDF load
 * Exoskeleton_sim_pybullet.py:14-18). Fails unless `device` is an sm_100 GPU. */
int exo_create(const ExoConfig* cfg, const ExoGeometry* geo, int32_t n_envs, int32_t device, ExoHandle** out);
int exo_destroy(ExoHandle* h);

/* read_env_texts result (Utilities/read_txt_env.py:73-170), columns 0-4 only:
 * tables[m][r][k], r = elbow_y, elbow_z, shoulder_x, shoulder_y, shoulder_z in degrees, padded to max_len. Host pointers. */
int exo_load_motions(ExoHandle* h, const double* tables, const int32_t* lengths, int32_t n_motions, int32_t max_len);

/* EXO_MOTOR_FULL only: link inertial data + solver settings (model == NULL: pybullet's defaults). Host pointers.
 * Call after exo_create, before exo_seed. */
int exo_set_motor_model(ExoHandle* h, const ExoLinkDynamics* links, const ExoMotorModel* model);
/* pybullet's defaults, for callers that want to change one knob */
void exo_default_motor_model(ExoMotorModel* out);

/* Per-env constructor arguments: tremor_sequence[7] (0/1) and reference_motion_file_num. Host pointers. */
int exo_set_env_types(ExoHandle* h, const uint8_t* tremor_sequence /*[n,7]*/, const int32_t* motion_id /*[n]*/);

/* np.random.seed(seeds[i]) for env i's private legacy-MT19937 stream (Utilities/seed_setting_.py:9), then,
 * if cfg.reference_ctor, the constructor's draws. Host pointer. Must follow exo_load_motions + exo_set_env_types. */
int exo_seed(ExoHandle* h, const uint32_t* seeds /*[n]*/, void* stream);

/* env.reset() (Exoskeleton_env.py:473-478 -> initialize_movement :193-254) for the envs whose mask byte is
 * non-zero (mask_dev == NULL: all). Writes obs rows of those envs only. */
int exo_reset(ExoHandle* h, const uint8_t* mask_dev, float* obs_dev /*[n,80]*/, void* stream);

/* env.step(action) (Exoskeleton_env.py:368-471) for every env. reward_dev is real [n]. */
int exo_step(ExoHandle* h, const float* actions_dev /*[n,7]*/, float* obs_dev /*[n,80]*/, void* reward_dev,
             uint8_t* done_dev /*[n]*/, const ExoInfo* info /*NULL: off*/, void* stream);

/* Same step with HOST buffers (pinned for full speed): H2D of actions, step, D2H of obs/reward/done, stream sync.
 * This is the call the single-env facade and the end-to-end benchmark use. From 16 384 envs up the population is
 * cut into 4 chunks (environment variable EXO_HOST_CHUNKS=1..8 overrides) whose copies and kernels overlap on
 * internal streams; results are identical to exo_step. */
int exo_step_host(ExoHandle* h, const float* actions_host, float* obs_host, void* reward_host, uint8_t* done_host,
                  void* stream);
int exo_reset_host(ExoHandle* h, const uint8_t* mask_host, float* obs_host, void* stream);

/* k = elements per env, dtype = EXO_DT_* of a field (valid after exo_set_env_types). */
int exo_field_info(ExoHandle* h, int32_t field, int32_t* k, int32_t* dtype);
/* resident state -> out_dev [n_envs, k] (env-major), and back. Asynchronous on `stream`. */
int exo_get_field(ExoHandle* h, int32_t field, void* out_dev, void* stream);
int exo_set_field(ExoHandle* h, int32_t field, const void* in_dev, void* stream);

/* Episode statistics accumulated on device since the last call with reset != 0:
 * out_dev[0..EXO_N_STATS) doubles: 0 episodes, 1 sum return, 2 sum length, 3 env-steps, 4 joint-limit crossings,
 * 5 steps with |torque|<|tremor| on >=1 of the 4 arm axes, 6 sum amplitude metric over steps where it is < 0,
 * 7 count of those steps; with `info` requested also the training script's suppression metrics
 * (Simulation/Exoskeleton_agent_train.py:150-193): 8 / 9 arm-axis samples with torque reduction >= 0 / < 0 (tremor_when_reduction),
 * 10 / 11 sum and count of negative torque-reduction percentages over the 7 axes, 12 / 13 the same for the amplitude
 * reduction (positive values dropped as by the script's `disregard`); EXO_MOTOR_FULL: 14 constraint-solver sweeps,
 * 15 env-steps through the general multibody step (14 / 15 = sweeps per step).
 * This vector is what ranks all-reduce over NCCL. */
int exo_stats(ExoHandle* h, double* out_dev, int32_t reset, void* stream);

/* Make `stream` wait (device side, no host sync) for every episode-parameter generation pass queued so far on the
 * handle's side stream. Not needed for correctness of exo_step / exo_reset / exo_get_field (they order themselves);
 * for timing the generation work and for callers that read handle memory by other means. */
int exo_wait_prefetch(ExoHandle* h, void* stream);

/* Debug/profiling: per-warp phase timestamps of the following exo_step launches are written to buf_dev, int64
 * [ceil(n_envs/32)][8] (NULL switches it off): globaltimer ns at entry, SM id, then clock64 at entry / all loads issued /
 * staged data landed / kinematics + actuators done / reward done / observation rows written. tools/step_timeline.py. */
int exo_debug_timeline(ExoHandle* h, void* buf_dev);

/* Toggle ExoConfig.auto_reset after creation (e.g. to time the step kernel alone). */
int exo_set_auto_reset(ExoHandle* h, int32_t on);

/* Number of kernels this handle has launched so far (for bench.py's gpu_launches). */
int64_t exo_launch_count(const ExoHandle* h);

/* Standalone batched forward_kinematics (Utilities/calculate_arm_end_effector_points.py:18-50):
 * theta_dev real [n,7] rad -> p_dev real [n,3]. */
int exo_forward_kinematics(ExoHandle* h, const void* theta_dev, void* p_dev, int32_t n, double L1, double L2, double L3,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif
