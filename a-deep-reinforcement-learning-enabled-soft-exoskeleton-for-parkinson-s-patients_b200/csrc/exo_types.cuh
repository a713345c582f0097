// exo_types.cuh — device-side layouts shared by the kernels of libexo_env.so.
//
// HBM layout: every per-env quantity is a field-major structure-of-arrays, element (field f, env e) at
// base[f * stride + e] with stride = n_envs rounded up to 32, so a warp touching one field of 32
// consecutive envs issues one fully coalesced 128 B (fp32) / 256 B (fp64) transaction.
#pragma once
#include <stdint.h>

#define EXO_MAX_LEN 512          // >= longest reference motion (347 rows)
#define EXO_SIGN_WORDS (EXO_MAX_LEN / 32)
#define EXO_STEP_THREADS 128

template <typename R>
struct EnvArrays {
    int n, stride;
    // carried state
    R* q;          // [5]  Bullet joint positions
    R* qlag;       // [5]  joint positions behind the link-state cache
    R* posvec;     // [21]
    R* elref;      // [3]
    R* fprev;      // [7]
    R* fprev2;     // [7]
    R* ret;        // [1] running episode return
    int* counts;   // [1]
    // per-episode parameters
    R* fmax;       // [2] shoulder, elbow
    R* shift;      // [42]
    R* G;          // [49]
    R* tt;         // [table_len][NAX]
    double* IDS;   // [147]
    double* mag;   // [1]
    // per-env constants
    int* meta;     // bits 0-6 tremor_sequence, bits 8-15 motion id
    uint32_t* mt;  // [625]
    // reset scratch (fp64)
    double* sc_noise;  // [max_len]
    double* sc_w1;     // [max_len]
    double* sc_w2;     // [max_len]
    uint32_t* sc_sign; // [7][EXO_SIGN_WORDS]
    // shared tables
    const R* motion;         // [n_motions][5][max_len] degrees
    const int* motion_len;   // [n_motions]
    int max_len, n_motions, table_len;
    double* stats;           // [EXO_N_STATS]
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
};

template <typename R>
struct StepConst {
    R fmax_sh0, fmax_el0;          // un-randomised maxima (obs normalisation, Exoskeleton_env.py:488-501)
    R L1, L2;                      // humerus / forearm length for the amplitude metric
    int lag, limits, auto_reset, metrics;
};

struct ResetConst {
    double amp_lo, amp_hi, f1_lo, f1_hi, f2_lo, f2_hi;
    double shift_range, act_range, noise_frac;
    double fmax_sh0, fmax_el0;
    double dt;
    int lag;
};

template <typename R>
struct InfoPtrs {
    R *actuator_torques, *torque_val, *ampl_val, *tremor_torque_val, *tremor_ampl_val, *reward_parts, *amp_metric;
};
