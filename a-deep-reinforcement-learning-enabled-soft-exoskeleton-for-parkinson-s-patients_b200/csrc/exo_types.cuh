// exo_types.cuh — device-side layouts shared by the kernels of libexo_env.so.
//
// HBM layout. Two access patterns meet in this library:
//   * exo_step_kernel   one THREAD per env, every control step: wants field-major arrays so that a warp touching
//                       one field of 32 consecutive envs issues one contiguous transaction;
//   * exo_reset_kernel  one WARP per env, once per episode: wants an env's private data contiguous.
// Everything the step touches is therefore stored as "vector-major SoA": the per-env record is cut into
// 4-word vectors V4<R> (16 B fp32 / 32 B fp64) and vector v of env e lives at base[v * stride + e]
// (stride = n_envs rounded up to 32). A warp loads vector v of its 32 envs as ONE 512 B (fp32) contiguous
// request made of 128-bit loads: a quarter of the load/store instructions of a word-major SoA and 4x the bytes
// in flight per instruction, which is what a latency-bound single-wave kernel needs. Data only the reset
// touches (MT19937 key, the fp64 I/D/S matrices) is env-major so the warp streams it coalesced.
#pragma once
#include <stdint.h>

#define EXO_MAX_LEN 512          // >= longest reference motion (347 rows)

#ifndef EXO_STEP_THREADS
#define EXO_STEP_THREADS 64      // 2 warps per block: 65 536 envs -> 1024 blocks ~ 6.9 per SM (balanced single wave)
#define EXO_STEP_MIN_BLOCKS 7    // <= 144 registers per thread: the 7 blocks (14 warps) an SM owns at 65 536 envs stay resident
#endif
#define EXO_OBS_PITCH 84         // floats per staged observation row: 21 16-byte chunks, odd -> conflict-free STS.128

// ---- word offsets inside the per-env records -------------------------------------------------------------
// carried state, 48 words = 12 vectors, rewritten every step
#define W_Q 0        // [5]  Bullet joint positions (rad): shoulder z,y,x, elbow y,z
#define W_QLAG 5     // [5]  joint positions the link-state cache was computed from
#define W_ELREF 10   // [3]  act_elbow_reference_positions as last read
#define W_POSVEC 13  // [21] position_vectors
#define W_FPREV 34   // [7]  prev_action (N)
#define W_FPREV2 41  // [7]  second_prev_action (N)
#define NV_STATE 12
// per-episode parameters, 44 words = 11 vectors, written at reset
#define W_SHIFT 0    // [42] dummy_shift_coordinates
#define W_FMAX 42    // [2]  max_output_shoulder, max_output_elbow (domain-randomised)
#define NV_EP 11
// propagator G (7x7) reordered so that the 4x4 block the plain step needs is vectors 0..3
#define NV_G 13
// general multibody step (ExoConfig.motor_model = EXO_MOTOR_FULL): 14 coupled degrees of freedom = 5 revolute arm joints
// + the 7 prismatic dummies riding on the humerus + the 2 on the forearm (the 5 dummies on the fixed base are decoupled
// 1-DoF systems their velocity motors hold at exactly 0). Per env: qd[14] | qp[9] prismatic positions | qplag[9] the
// positions behind the link-state cache = 32 words = 8 vectors; the step kernel hands the position targets over in
// 2 more vectors: tgt[5], live flag.
#define EXO_NDOF 14
#define EXO_NPRIS 9
#define W_QD 0
#define W_QP 14
#define W_QPLAG 23
#define NV_DYN 8
#define NV_TGT 2
#define EXO_MAX_ROWS 24          // 14 motor rows + 2 limit rows for each of the 5 revolute joints
// Per-episode parameters are DOUBLE-BUFFERED: while an env plays the episode held in slot s, the parameters of its next
// episode are generated into slot s^1 on a side stream (exo_gen_kernel), one episode ahead, so the end of an episode is a
// slot flip + first observation inside the step kernel and nothing waits for the ~50 us generation chain. The slot lives
// in bit 16 of `counts` (no extra load). A ring of EXO_RING work lists carries the envs whose spare slot needs refilling.
#define EXO_SLOT_SHIFT 16
#define EXO_COUNT_MASK 0xffff
#define EXO_RING 32
#define EXO_MT_PITCH 640         // words per env in the MT19937 array: 624 key + position (+pad to 128 B multiple)
#define EXO_IDS_PITCH 148        // doubles per env: I, D, S (3 x 49) + pad

__host__ __device__ __forceinline__ int exo_g_word(int r, int c) {
    return (r < 4) ? (c < 4 ? r * 4 + c : 16 + r * 3 + (c - 4)) : 28 + (r - 4) * 7 + c;
}

template <typename R>
struct __align__(16) V4 { R v[4]; };

// scalar word w of env e inside a vector-major array
template <typename R>
__host__ __device__ __forceinline__ R* exo_word(V4<R>* base, int w, int e, int stride) {
    return &base[(size_t)(w >> 2) * stride + e].v[w & 3];
}

template <typename R>
struct EnvArrays {
    int n, stride;
    V4<R>* state;    // [NV_STATE][stride]
    V4<R>* ep;       // [2][NV_EP][stride]                  slot-major: slot s at + s * ep_slot
    V4<R>* G;        // [2][NV_G][stride]
    V4<R>* tt;       // [2][table_len][stride][NAXP/4]      tremor table, time-major: one step reads rows c-1, c, c+1
    size_t ep_slot, G_slot, tt_slot;                        // slot strides, in V4 units
    R* ret;          // [stride] running episode return
    int* counts;     // [stride] self.counts | slot << 16
    int* meta;       // [stride] bits 0-6 tremor_sequence, bits 8-15 motion id, bits 16-31 motion length (max_count)
    double* IDS;     // [2][n][EXO_IDS_PITCH]  I_current, D_current, S_current (fp64, env-major), slot s at + s * n * EXO_IDS_PITCH
    double* mag;     // [2][stride] tremor_magnitude
    uint32_t* mt;    // [n][EXO_MT_PITCH]   numpy legacy MT19937 key + position (env-major)
    const R* motion;         // [n_motions][5][max_len] degrees
    const int* motion_len;   // [n_motions]
    int max_len, n_motions, table_len;
    V4<R>* dyn;              // [NV_DYN][stride]  EXO_MOTOR_FULL only: qd, qp, qplag
    V4<R>* tgt;              // [NV_TGT][stride]  EXO_MOTOR_FULL only: this step's position targets + live flag
    double* stats;           // [EXO_N_STATS]
    int* fin_count;          // [EXO_RING]          envs whose episode began in call t (ring slot t mod EXO_RING): their spare
    int* fin_list;           // [EXO_RING][stride]  slot is refilled by exo_gen_kernel on the side stream
};

// Geometry of the 5-revolute chain, flattened on the host from the URDF tree (exo_api.cu: build_chain).
template <typename R>
struct ChainConst {
    R Rorg[5][9];     // joint-origin rotation of revolute link i in its parent frame
    R porg[5][3];     // joint-origin translation
    R p0[3];          // world position of link-0 frame origin
    R off_hum[7][3];  // act11, act21, act32, act42, act52, act62, act72 in the humerus (link 2) frame
    R off_elbow[3];   // elbow reference (auxlink3 COM) in the link-3 frame
    R off_alk[2][3];  // act12, act22 in the forearm (link 4) frame
    R fixed[5][3];    // act31, act41, act51, act61, act71 world positions (base-fixed)
    R sh_ref[3];      // shoulder reference (link-0 COM) world position
    R lower[5], upper[5];
    R pax_hum[7][3];  // prismatic axis of each humerus dummy in the humerus frame (EXO_MOTOR_FULL: endpoint += axis * qp)
    R pax_alk[2][3];  // same for the two forearm dummies
};

// Inertial data and solver settings of the general multibody step (exo_motor.cuh), flattened on the host from the
// URDF link data (exo_api.cu: build_dynamics). DoF order: revolute 0..4, humerus dummies in off_hum order, forearm dummies.
template <typename R>
struct DynConst {
    R m_link[5], com_link[5][3], I_link[5][6];     // link frame: mass, COM, inertia tensor (xx,xy,xz,yy,yz,zz)
    R m_d[EXO_NPRIS], I_d[EXO_NPRIS][6];           // dummy mass, inertia tensor in the carrier frame
    R kp, kd, cap_pos, cap_vel;                    // position motors (joints 0-4) / default velocity motors (dummies)
    R k_lin, k_ang, g[3], dt, erp, limit_cap, threshold;
    int iterations, n_cons;
    signed char cons[EXO_MAX_ROWS];                // solver order: dof (motor row) or -(dof+1) (joint-limit constraint)
    signed char slot_dof[EXO_NDOF], dof_slot[EXO_NDOF];   // solver slot r = r-th motor row of `cons` <-> physical DoF
    unsigned char tri[EXO_NDOF][EXO_NDOF];         // index of (a, b) = (b, a) among the 105 stored entries of the symmetric M^-1
    unsigned char etri[19][19];                    // tri[slot of the joint of entry a][slot of the joint of entry b], entries of `cons`
};

template <typename R>
struct StepConst {
    R fmax_sh0, fmax_el0;          // un-randomised maxima (obs normalisation, Exoskeleton_env.py:488-501)
    R inv_fmax_sh0, inv_fmax_el0;
    R L1, L2;                      // humerus / forearm length for the amplitude metric
    int lag, limits, auto_reset, metrics, motor_full, ode_rk45;
    long long* timeline;           // debug (EXO_STEP_TIMELINE): per warp 8 words: globaltimer at entry, smid, clock64 at 6 phase marks
    int stagger_cycles, n_sm;      // blocks of dispatch layer l = blockIdx / n_sm start their loads l * stagger_cycles late
    int list;                      // ring slot of the work list this step appends to (it clears the next one)
    int fused_begin;               // auto-reset: finished envs flip to their prefetched episode inside this kernel
    int env0, env1;                // env range [env0, env1) of this launch (env0 a multiple of 32): exo_step_host pipelines
                                   // chunks of the population against their PCIe copies
};

struct ResetConst {
    double amp_lo, amp_hi, f1_lo, f1_hi, f2_lo, f2_hi;
    double shift_range, act_range, noise_frac;
    double fmax_sh0, fmax_el0;
    double dt;
    int lag;
    int motor_full;    // EXO_MOTOR_FULL: the actuator end points follow their prismatic dummies
    int ctor_skip;     // 32-bit outputs consumed before the episode draws (2 for the constructor's rand(), env:70)
    int list;          // >= 0: the envs of work list `list` of the ring; < 0: by mask (NULL: all)
    int append;        // exo_begin_kernel, mask mode: append the envs it started to work list `append` (>= 0)
};

template <typename R>
struct InfoPtrs {
    R *actuator_torques, *torque_val, *ampl_val, *tremor_torque_val, *tremor_ampl_val, *reward_parts, *amp_metric;
};

// field ids of exo_get_field / exo_set_field (include/exo_env.h EXO_FIELD_*)
enum {
    F_Q = 0, F_Q_LAG, F_POS_VEC, F_ELBOW_REF, F_PREV_ACTION, F_PREV_ACTION2, F_FORCE_MAX, F_SHIFT, F_G, F_IDS,
    F_TREMOR_TABLE, F_TREMOR_MAG, F_COUNTS, F_META, F_MT_STATE, F_EP_RETURN, F_QD, F_Q_PRIS, F_Q_PRIS_LAG, F_N_FIELDS
};
