"""ctypes binding of libexo_env.so (include/exo_env.h). Loading fails loudly: there is no CPU path."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("EXO_LIB", os.path.join(HERE, "libexo_env.so"))   # EXO_LIB: tuning builds of the same sources

OBS_DIM, ACT_DIM, NLINKS, N_STATS = 80, 7, 19, 16
F32, F64 = 0, 1
ODE_PROPAGATOR, ODE_RK45 = 0, 1
MOTOR_IDEAL, MOTOR_FULL = 0, 1


class ExoConfig(C.Structure):
    _fields_ = [("tremor_amplitude_range", C.c_double * 2), ("first_harmonics_interval", C.c_double * 2),
                ("second_harmonics_interval", C.c_double * 2),
                ("max_force_shoulder", C.c_double), ("max_force_elbow", C.c_double),
                ("dr_actuator_end_pos_shift", C.c_double), ("dr_actuator_range", C.c_double),
                ("matrix_noise_fraction", C.c_double),
                ("humerus_length", C.c_double), ("forearm_length", C.c_double), ("hand_length", C.c_double),
                ("link_state_lag", C.c_int32), ("joint_limits", C.c_int32), ("precision", C.c_int32),
                ("ode", C.c_int32), ("auto_reset", C.c_int32), ("reference_ctor", C.c_int32),
                ("metrics", C.c_int32), ("motor_model", C.c_int32)]


class ExoGeometry(C.Structure):
    _fields_ = [("parent", C.c_int32 * NLINKS), ("joint_type", C.c_int32 * NLINKS),
                ("origin_xyz", (C.c_double * 3) * NLINKS), ("origin_R", (C.c_double * 9) * NLINKS),
                ("axis", (C.c_double * 3) * NLINKS), ("com_xyz", (C.c_double * 3) * NLINKS),
                ("lower", C.c_double * NLINKS), ("upper", C.c_double * NLINKS), ("base_pos", C.c_double * 3),
                ("act_link", C.c_int32 * 14), ("shoulder_link", C.c_int32), ("elbow_link", C.c_int32)]


class ExoLinkDynamics(C.Structure):
    _fields_ = [("mass", C.c_double * NLINKS), ("inertia_diag", (C.c_double * 3) * NLINKS),
                ("inertial_R", (C.c_double * 9) * NLINKS), ("box_size", (C.c_double * 3) * NLINKS),
                ("box_xyz", (C.c_double * 3) * NLINKS), ("box_R", (C.c_double * 9) * NLINKS)]


class ExoMotorModel(C.Structure):
    _fields_ = [("inertia_from_file", C.c_int32), ("iterations", C.c_int32), ("row_order", C.c_int32), ("reserved", C.c_int32),
                ("residual_threshold", C.c_double), ("kp", C.c_double), ("kd", C.c_double), ("max_force", C.c_double),
                ("default_motor_impulse", C.c_double), ("linear_damping", C.c_double), ("angular_damping", C.c_double),
                ("gravity", C.c_double * 3), ("erp", C.c_double), ("limit_max_impulse", C.c_double)]


class ExoPolicyWeights(C.Structure):
    """include/exo_policy.h: host pointers to torch.nn.Linear parameters (weight [out, in] row-major float32, bias [out])."""
    NAMES = ("zs1", "zs2", "zs3", "l0", "l1", "l2", "l3")
    _fields_ = [(n + s, C.POINTER(C.c_float)) for n in NAMES for s in ("_w", "_b")]


class ExoTD7Config(C.Structure):
    """include/exo_td7.h: Hyperparameters + network sizes of the TD7 learner."""
    _fields_ = [(k, C.c_int32) for k in ("state_dim", "action_dim", "zs_dim", "enc_hdim", "critic_hdim", "actor_hdim", "batch_rows",
                                         "target_update_rate", "policy_freq")] + \
               [(k, C.c_float) for k in ("discount", "target_policy_noise", "noise_clip", "policy_noise_decrease", "alpha",
                                         "min_priority", "actor_lr", "critic_lr", "encoder_lr", "weight_decay")] + \
               [("seed", C.c_uint64), ("offline", C.c_int32), ("lmbda", C.c_float)]


EXO_TD7_GRAD_HOOK = C.CFUNCTYPE(None, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p)


class ExoInfo(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("actuator_torques", "torque_val", "ampl_val", "tremor_torque_val",
                                          "tremor_ampl_val", "reward_parts", "amp_metric")]


# EXO_FIELD_* (include/exo_env.h): name -> id; EXO_DT_* -> torch dtype name
FIELDS = {name: i for i, name in enumerate(
    ("q", "q_lag", "pos_vec", "elbow_ref", "prev_action", "prev_action2", "force_max", "shift", "G", "IDS", "tremor_table",
     "tremor_mag", "counts", "meta", "mt_state", "ep_return", "qd", "q_pris", "q_pris_lag"))}
FIELD_NEXT = 0x100        # EXO_FIELD_NEXT: the prefetched copy (next episode's parameters) of a per-episode field
DT_REAL, DT_F64, DT_I32, DT_U32 = 0, 1, 2, 3


# name -> (restype, argtypes); also the list tests check against include/exo_env.h
SIGNATURES = {
    "exo_last_error": (C.c_char_p, []),
    "exo_abi_version": (C.c_int, []),
    "exo_create": (C.c_int, [C.POINTER(ExoConfig), C.POINTER(ExoGeometry), C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "exo_destroy": (C.c_int, [C.c_void_p]),
    "exo_set_motor_model": (C.c_int, [C.c_void_p, C.POINTER(ExoLinkDynamics), C.POINTER(ExoMotorModel)]),
    "exo_default_motor_model": (None, [C.POINTER(ExoMotorModel)]),
    "exo_load_motions": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32), C.c_int32, C.c_int32]),
    "exo_set_env_types": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint8), C.POINTER(C.c_int32)]),
    "exo_seed": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p]),
    "exo_reset": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(ExoInfo), C.c_void_p]),
    "exo_step_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_reset_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_field_info": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "exo_get_field": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "exo_set_field": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "exo_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "exo_wait_prefetch": (C.c_int, [C.c_void_p, C.c_void_p]),
    "exo_debug_timeline": (C.c_int, [C.c_void_p, C.c_void_p]),
    "exo_set_auto_reset": (C.c_int, [C.c_void_p, C.c_int32]),
    "exo_launch_count": (C.c_int64, [C.c_void_p]),
    "exo_forward_kinematics": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_void_p]),
    # include/exo_lap.h
    "exo_lap_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_double, C.c_int32, C.c_int32,
                                 C.POINTER(C.c_void_p)]),
    "exo_lap_destroy": (C.c_int, [C.c_void_p]),
    "exo_lap_add": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_int32, C.c_void_p]),
    "exo_lap_sample": (C.c_int, [C.c_void_p] * 8 + [C.c_void_p]),
    "exo_lap_update_priority": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_lap_reset_max_priority": (C.c_int, [C.c_void_p, C.c_void_p]),
    "exo_lap_reset_buffer": (C.c_int, [C.c_void_p, C.c_void_p]),
    "exo_lap_counters": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                   C.POINTER(C.c_double), C.c_void_p]),
    "exo_lap_storage": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    "exo_lap_launch_count": (C.c_int64, [C.c_void_p]),
    # include/exo_policy.h
    "exo_policy_create": (C.c_int, [C.c_int32, C.POINTER(C.c_void_p)]),
    "exo_policy_destroy": (C.c_int, [C.c_void_p]),
    "exo_policy_set_weights": (C.c_int, [C.c_void_p, C.POINTER(ExoPolicyWeights), C.c_int32, C.c_int32, C.c_int32]),
    "exo_policy_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_policy_forward_fp32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_policy_forward_fp32_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_policy_profile": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(C.c_int64), C.c_void_p]),
    "exo_policy_launch_count": (C.c_int64, [C.c_void_p]),
    # include/exo_td7.h
    "exo_td7_create": (C.c_int, [C.c_int32, C.POINTER(ExoTD7Config), C.POINTER(C.c_void_p)]),
    "exo_td7_destroy": (C.c_int, [C.c_void_p]),
    "exo_td7_net_size": (C.c_int64, [C.c_void_p, C.c_int32]),
    "exo_td7_set_net": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "exo_td7_get_net": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "exo_td7_copy_net": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "exo_td7_set_opt": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_int64]),
    "exo_td7_get_opt": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "exo_td7_train": (C.c_int, [C.c_void_p] + [C.c_void_p] * 7 + [C.POINTER(C.c_int32), C.c_void_p]),
    "exo_td7_scalars": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "exo_td7_set_scalars": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "exo_td7_set_grad_hook": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "exo_td7_randn": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "exo_td7_launch_count": (C.c_int64, [C.c_void_p]),
}

_lib = None


class ExoError(RuntimeError):
    pass


def load():
    """dlopen libexo_env.so and bind every entry point. Raises if the library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ExoError("%s is missing: run `python __graft_entry__.py build` (nvcc, sm_100a). "
                           "There is no CPU implementation to fall back to." % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise ExoError("libexo_env: %s (status %d)" % (load().exo_last_error().decode(), rc))
