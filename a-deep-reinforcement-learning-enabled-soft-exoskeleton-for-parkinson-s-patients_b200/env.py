"""Host-side mirror of the reference's environment interface, backed by libexo_env.so.

  BatchedExoskeletonEnv   N environments resident on one B200: reset() / step(actions[N,7]) -> (obs, reward, done)
  ExoskeletonEnv_train    single-env facade with the reference's constructor and return types
                          (Environment/Exoskeleton_env.py:34-605), so Simulation/Exoskeleton_agent_train.py's loop,
                          Agent.select_action and LAP.add consume it unchanged.

PyTorch is used for device memory and streams only; all env arithmetic runs in the CUDA kernels.
"""
import ctypes as C

import numpy as np
import torch

from . import _capi, motions as _motions, urdf as _urdf

INFO_KEYS = ("actuator_torques", "torque_val", "ampl_val", "tremor_torque_val", "tremor_ampl_val")
REWARD_KEYS = ("reward_unwanted", "reward_torque", "reward_axis", "reward_control", "reward_smoothness")
STAT_NAMES = ("episodes", "return_sum", "length_sum", "env_steps", "limit_crossings", "steps_with_arm_reduction",
              "amp_metric_neg_sum", "amp_metric_neg_count", "arm_axis_no_reduction_count", "arm_axis_reduction_count",
              "torque_reduction_neg_sum", "torque_reduction_neg_count", "ampl_reduction_neg_sum", "ampl_reduction_neg_count")


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class BatchedExoskeletonEnv(object):
    """N independent ExoskeletonEnv_train instances stepped by one fused kernel launch.

    motion_ids[i] / tremor_sequences[i] / seeds[i] are env i's `reference_motion_file_num`, `tremor_sequence`
    and the argument of the `np.random.seed` call that precedes its construction in the reference
    (Utilities/seed_setting_.py:9). Every env owns a private numpy-legacy MT19937 stream on the device, so env i
    reproduces, draw for draw, a reference env built right after np.random.seed(seeds[i]).
    """

    def __init__(self, motion_ids, tremor_sequences, seeds=None, *, tremor_amplitude_range, first_harmonics_interval,
                 second_harmonics_interval, max_force_shoulder, max_force_elbow, dr_actuator_end_pos_shift=0.02,
                 dr_actuator_range=0.1, matrix_noise_fraction=0.1, device="cuda:0", precision="fp32", auto_reset=False,
                 reference_ctor=True, link_state_lag=1, joint_limits="off", metrics=False, motions=None, urdf=None,
                 humerus_length=0.4, forearm_length=0.4, hand_length=0.05, motor_model="ideal", motor_options=None,
                 ode="propagator"):
        """motor_model: "ideal" - the fixed point of Bullet's motor solve, q+ = q + kp (target - q) (default);
        "full" - the general multibody step (forward dynamics + 50-sweep projected Gauss-Seidel, csrc/exo_motor.cuh);
        motor_options: dict of ExoMotorModel fields overriding pybullet's defaults (inertia_from_file, iterations,
        row_order, residual_threshold, kp, kd, max_force, default_motor_impulse, linear_damping, angular_damping, ...)."""
        self.lib = _capi.load()
        if not torch.cuda.is_available():
            raise _capi.ExoError("no CUDA device: the batched env runs on a B200 only (there is no CPU fallback)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _capi.ExoError("device must be a CUDA device, got %s" % device)
        motion_ids = np.ascontiguousarray(motion_ids, dtype=np.int32).reshape(-1)
        self.n = int(motion_ids.size)
        seqs = np.ascontiguousarray(tremor_sequences, dtype=np.uint8)
        if seqs.ndim == 1:
            seqs = np.broadcast_to(seqs, (self.n, 7)).copy()
        if seqs.shape != (self.n, 7):
            raise ValueError("tremor_sequences must be [N,7] or [7]")
        if seeds is None:
            seeds = np.arange(self.n, dtype=np.uint32)
        seeds = np.ascontiguousarray(np.asarray(seeds, dtype=np.int64) & 0xFFFFFFFF, dtype=np.uint32).reshape(-1)
        if seeds.size != self.n:
            raise ValueError("seeds must have one entry per env")
        self.motion_ids, self.tremor_sequences, self.seeds = motion_ids, seqs, seeds
        self.precision = {"fp32": _capi.F32, "fp64": _capi.F64}[precision]
        self.real = torch.float32 if self.precision == _capi.F32 else torch.float64
        self.auto_reset = bool(auto_reset)

        cfg = _capi.ExoConfig()
        cfg.tremor_amplitude_range[:] = [float(x) for x in tremor_amplitude_range]
        cfg.first_harmonics_interval[:] = [float(x) for x in first_harmonics_interval]
        cfg.second_harmonics_interval[:] = [float(x) for x in second_harmonics_interval]
        cfg.max_force_shoulder, cfg.max_force_elbow = float(max_force_shoulder), float(max_force_elbow)
        cfg.dr_actuator_end_pos_shift = float(dr_actuator_end_pos_shift)
        cfg.dr_actuator_range = float(dr_actuator_range)
        cfg.matrix_noise_fraction = float(matrix_noise_fraction)
        cfg.humerus_length, cfg.forearm_length, cfg.hand_length = float(humerus_length), float(forearm_length), float(hand_length)
        cfg.link_state_lag = int(link_state_lag)
        cfg.joint_limits = {"off": 0, "clamp": 1}[joint_limits]
        cfg.precision = self.precision
        cfg.ode = {"propagator": _capi.ODE_PROPAGATOR, "rk45": _capi.ODE_RK45}[ode]     # "rk45": fp64 verification mode (exo_rk45.cuh)
        cfg.auto_reset = int(self.auto_reset)
        cfg.reference_ctor = int(bool(reference_ctor))
        cfg.metrics = int(bool(metrics))
        cfg.motor_model = {"ideal": _capi.MOTOR_IDEAL, "full": _capi.MOTOR_FULL}[motor_model]
        self.motor_model = motor_model
        self.cfg = cfg
        self.max_force_shoulder, self.max_force_elbow = float(max_force_shoulder), float(max_force_elbow)

        self.geometry = _urdf.load_geometry(urdf)
        geo = _urdf.to_struct(self.geometry)
        self.motions = _motions.load_motions(motions)
        tables, lengths = _motions.pack(self.motions)
        self.motion_lengths = lengths
        self.max_count = lengths[motion_ids]                            # per env (Exoskeleton_env.py:63)
        self.max_reward = seqs.sum(1) * 0.5 + 0.9 + 0.05 + 0.05 + 0.5   # Exoskeleton_env.py:167-169

        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", dev_index)
        self.h = C.c_void_p()
        with torch.cuda.device(dev_index):
            _capi.check(self.lib.exo_create(C.byref(cfg), C.byref(geo), self.n, dev_index, C.byref(self.h)))
            if motor_model == "full":
                mm = _capi.ExoMotorModel()
                self.lib.exo_default_motor_model(C.byref(mm))
                for k, v in (motor_options or {}).items():
                    if k == "gravity":
                        mm.gravity[:] = [float(x) for x in v]
                    else:
                        setattr(mm, k, v)
                links = _urdf.to_dynamics_struct(self.geometry)
                _capi.check(self.lib.exo_set_motor_model(self.h, C.byref(links), C.byref(mm)))
            _capi.check(self.lib.exo_load_motions(self.h, tables.ctypes.data_as(C.POINTER(C.c_double)),
                                                  lengths.ctypes.data_as(C.POINTER(C.c_int32)), len(self.motions),
                                                  tables.shape[2]))
            _capi.check(self.lib.exo_set_env_types(self.h, seqs.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                   motion_ids.ctypes.data_as(C.POINTER(C.c_int32))))
            _capi.check(self.lib.exo_seed(self.h, seeds.ctypes.data_as(C.POINTER(C.c_uint32)), self._stream()))
        dev = self.device
        self._obs = [torch.zeros(self.n, 80, dtype=torch.float32, device=dev) for _ in range(2)]
        self._flip = 0
        self._reward = torch.zeros(self.n, dtype=self.real, device=dev)
        self._done = torch.zeros(self.n, dtype=torch.uint8, device=dev)
        self._done_bool = self._done.view(torch.bool)      # the kernel writes 0/1 bytes: a view, not a conversion kernel
        self._info = None
        self._stats = torch.zeros(_capi.N_STATS, dtype=torch.float64, device=dev)
        self._host = None

    # ------------------------------------------------------------------ plumbing
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.exo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def wait_prefetch(self):
        """The current stream waits for the episode-parameter generation passes queued so far (side stream); for timing."""
        _capi.check(self.lib.exo_wait_prefetch(self.h, self._stream()))

    def set_auto_reset(self, on):
        _capi.check(self.lib.exo_set_auto_reset(self.h, int(bool(on))))
        self.auto_reset = bool(on)

    @property
    def launch_count(self):
        return int(self.lib.exo_launch_count(self.h))

    # ------------------------------------------------------------------ reference-shaped API, vectorised
    def reset(self, mask=None):
        """env.reset() for all envs (mask=None) or those with mask[i] != 0. Returns obs[N,80] (float32, device)."""
        obs = self._obs[self._flip]
        m = None
        if mask is not None:
            m = torch.as_tensor(mask, device=self.device).to(torch.uint8).contiguous()
        _capi.check(self.lib.exo_reset(self.h, _ptr(m), _ptr(obs), self._stream()))
        return obs

    def step(self, actions, with_info=False, keep_idle_obs=True):
        """env.step for every env. actions: float32 [N,7] on the device. Returns (obs, reward, done) and, with
        with_info, the reference's info dict as [N,7] / [N] tensors. The obs tensor returned by the previous call
        stays valid (double buffered) so a replay buffer can store (obs, action, next_obs). Without auto_reset,
        envs that already reported done idle; keep_idle_obs carries their last observation into the new buffer."""
        a = torch.as_tensor(actions, device=self.device, dtype=torch.float32).contiguous()
        if a.shape != (self.n, 7):
            raise ValueError("actions must be [N,7]")
        self._flip ^= 1
        obs = self._obs[self._flip]
        if keep_idle_obs and not self.auto_reset:
            obs.copy_(self._obs[self._flip ^ 1])     # idle (finished) envs keep their last observation
        info_struct = None
        if with_info:
            if self._info is None:
                self._info = {k: torch.zeros(self.n, 7, dtype=self.real, device=self.device) for k in INFO_KEYS}
                self._info["reward_parts"] = torch.zeros(self.n, 5, dtype=self.real, device=self.device)
                self._info["amp_metric"] = torch.zeros(self.n, dtype=self.real, device=self.device)
            s = _capi.ExoInfo()
            for k in INFO_KEYS + ("reward_parts", "amp_metric"):
                setattr(s, k, self._info[k].data_ptr())
            info_struct = C.byref(s)
        _capi.check(self.lib.exo_step(self.h, _ptr(a), _ptr(obs), _ptr(self._reward), _ptr(self._done), info_struct,
                                      self._stream()))
        if with_info:
            return obs, self._reward, self._done_bool, self._info
        return obs, self._reward, self._done_bool

    def _host_buffers(self):
        if self._host is None:
            self._host = dict(obs=torch.zeros(self.n, 80, dtype=torch.float32).pin_memory(),
                              reward=torch.zeros(self.n, dtype=self.real).pin_memory(),
                              done=torch.zeros(self.n, dtype=torch.uint8).pin_memory())
        return self._host

    def step_host(self, actions_host):
        """The same step through HOST buffers (exo_step_host): H2D actions, kernels, D2H obs/reward/done, sync.
        actions_host: contiguous float32 CPU tensor [N,7] (pinned for full speed). Returns pinned CPU tensors
        (reused between calls)."""
        o = self._host_buffers()
        if actions_host.dtype != torch.float32 or actions_host.device.type != "cpu" or not actions_host.is_contiguous():
            raise ValueError("actions_host must be a contiguous float32 CPU tensor")
        _capi.check(self.lib.exo_step_host(self.h, _ptr(actions_host), _ptr(o["obs"]), _ptr(o["reward"]), _ptr(o["done"]),
                                           self._stream()))
        return o["obs"], o["reward"], o["done"]

    def reset_host(self, mask_host=None):
        o = self._host_buffers()
        m = None
        if mask_host is not None:
            m = torch.as_tensor(mask_host).to(torch.uint8).contiguous()
        _capi.check(self.lib.exo_reset_host(self.h, _ptr(m), _ptr(o["obs"]), self._stream()))
        return o["obs"]

    def stats(self, reset=False):
        """Device-side episode statistics, float64 [16] (see STAT_NAMES); the vector ranks all-reduce over NCCL."""
        _capi.check(self.lib.exo_stats(self.h, _ptr(self._stats), int(reset), self._stream()))
        return self._stats

    def forward_kinematics(self, theta, L1=0.4, L2=0.4, L3=0.05):
        th = torch.as_tensor(theta, device=self.device, dtype=self.real).contiguous().reshape(-1, 7)
        p = torch.empty(th.shape[0], 3, dtype=self.real, device=self.device)
        _capi.check(self.lib.exo_forward_kinematics(self.h, _ptr(th), _ptr(p), th.shape[0], L1, L2, L3, self._stream()))
        return p

    # ------------------------------------------------------------------ resident state by field (tests, checkpoints)
    def _field_info(self, name):
        k, dt = C.c_int32(), C.c_int32()
        _capi.check(self.lib.exo_field_info(self.h, _capi.FIELDS[name], C.byref(k), C.byref(dt)))
        dtype = {_capi.DT_REAL: self.real, _capi.DT_F64: torch.float64, _capi.DT_I32: torch.int32, _capi.DT_U32: torch.int32}[dt.value]
        return k.value, dtype

    PER_EPISODE_FIELDS = ("force_max", "shift", "G", "IDS", "tremor_mag", "tremor_table")

    def get_field(self, name, next=False):
        """Copy of one state field as an env-major tensor [N, k] (tremor_table: [N, table_len, axes]); the layout
        inside the handle is private (exo_get_field gathers it on the device). Episode parameters are generated one episode
        ahead: next=True reads the prefetched copy (what the env's NEXT episode will use) of a PER_EPISODE_FIELDS entry."""
        k, dtype = self._field_info(name)
        out = torch.empty(self.n, k, dtype=dtype, device=self.device)
        fid = _capi.FIELDS[name] | (_capi.FIELD_NEXT if next else 0)
        _capi.check(self.lib.exo_get_field(self.h, fid, _ptr(out), self._stream()))
        if name == "tremor_table":
            axes = 7 if int(self.tremor_sequences[:, 4:].max(initial=0)) else 4
            return out.reshape(self.n, k // axes, axes)
        return out

    def set_field(self, name, value, next=False):
        k, dtype = self._field_info(name)
        val = torch.as_tensor(value, device=self.device).to(dtype).reshape(self.n, k).contiguous()
        fid = _capi.FIELDS[name] | (_capi.FIELD_NEXT if next else 0)
        _capi.check(self.lib.exo_set_field(self.h, fid, _ptr(val), self._stream()))
        torch.cuda.current_stream(self.device).synchronize()      # `val` may be a temporary

    CHECKPOINT_FIELDS = ("q", "q_lag", "pos_vec", "elbow_ref", "prev_action", "prev_action2", "force_max", "shift", "G",
                         "ep_return", "IDS", "tremor_mag", "counts", "tremor_table", "mt_state")

    def state_dict(self):
        """Everything needed to resume: a handful of SoA tensors - the carried state, the running episode's parameters, the
        prefetched parameters of the next episode ("next/<field>") and the RNG stream position behind them."""
        sd = {k: self.get_field(k) for k in self.CHECKPOINT_FIELDS}
        sd.update({"next/" + k: self.get_field(k, next=True) for k in self.PER_EPISODE_FIELDS})
        if self.motor_model == "full":
            sd.update({k: self.get_field(k) for k in ("qd", "q_pris", "q_pris_lag")})
        return sd

    def load_state_dict(self, sd):
        for k in self.CHECKPOINT_FIELDS:
            self.set_field(k, sd[k])
        for k in self.PER_EPISODE_FIELDS:
            self.set_field(k, sd["next/" + k], next=True)
        if self.motor_model == "full":
            for k in ("qd", "q_pris", "q_pris_lag"):
                self.set_field(k, sd[k])


class _Box(object):
    """Stand-in for gym.spaces.Box (gym is optional): shape/low/high/dtype, which is all callers read
    (Exoskeleton_agent_train.py:80)."""

    def __init__(self, low, high, dtype=np.float32):
        self.low = np.asarray(low, dtype=dtype)
        self.high = np.asarray(high, dtype=dtype)
        self.shape = self.low.shape
        self.dtype = np.dtype(dtype)

    def sample(self):
        return np.random.uniform(self.low, self.high).astype(self.dtype)


class ExoskeletonEnv_train(object):
    """Drop-in for Environment.Exoskeleton_env.ExoskeletonEnv_train: same constructor arguments
    (Exoskeleton_env.py:38-48), reset() -> (obs[80] float32, 2) (:473-478), step(a[7]) -> (obs, reward: float,
    done: bool, False, info) (:368-471) with the same info keys, and the helper methods callers use.

    Difference that cannot be hidden: the reference draws episode parameters from numpy's *global* stream, shared
    with every other env and with the warm-up policy; here each env owns a private MT19937 stream on the device,
    seeded with `seed` (default: one draw from numpy's global stream, so seeding numpy upstream stays reproducible).
    """
    metadata = {"render.modes": ["human"]}

    def __init__(self, reference_motion_file_num, tremor_sequence, tremor_amplitude_range, first_harmonics_interval,
                 second_harmonics_interval, max_force_shoulder, max_force_elbow, dr_actuator_end_pos_shift=0.02,
                 dr_actuator_range=0.1, matrix_noise_fraction=0.1, *, seed=None, device="cuda:0", precision="fp64", **kw):
        if seed is None:
            seed = int(np.random.randint(0, 2 ** 31 - 1))
        self.file_num = str(reference_motion_file_num)
        self.tremor_input_sequence = np.asarray(tremor_sequence)
        self.batch = BatchedExoskeletonEnv([int(self.file_num)], [self.tremor_input_sequence.astype(np.uint8)], [seed],
                                           tremor_amplitude_range=tremor_amplitude_range,
                                           first_harmonics_interval=first_harmonics_interval,
                                           second_harmonics_interval=second_harmonics_interval,
                                           max_force_shoulder=max_force_shoulder, max_force_elbow=max_force_elbow,
                                           dr_actuator_end_pos_shift=dr_actuator_end_pos_shift,
                                           dr_actuator_range=dr_actuator_range, matrix_noise_fraction=matrix_noise_fraction,
                                           device=device, precision=precision, auto_reset=False, **kw)
        self.max_count = int(self.batch.max_count[0])
        self.max_reward = float(self.batch.max_reward[0])
        self.dt = 1 / 40
        self.counts = 2
        self.action_space = _Box(-np.ones(7), np.ones(7))
        low = np.concatenate([np.repeat(0.0, 14), np.repeat(-15.0, 12), np.repeat(-2.0, 54)])
        high = np.concatenate([np.repeat(1.5, 14), np.repeat(15.0, 12), np.repeat(2.0, 54)])
        self.observation_space = _Box(low, high)
        self.state = None
        self._out = None

    def reset(self):
        obs = self.batch.reset()
        self.counts = 2
        self.state = obs[0].cpu().numpy()
        return self.state, self.counts

    def _blob(self):
        """All outputs of one step in ONE device buffer (and one pinned host mirror), so a step costs one H2D copy of the
        action and one D2H copy instead of a synchronising copy per output: obs | reward | done | info[5x7] | reward parts[5] |
        amplitude metric, each 16-byte aligned."""
        if self._out is None:
            b = self.batch
            rb = 8 if b.real == torch.float64 else 4
            offs, o = {}, 0
            for name, nbytes in (("obs", 320), ("reward", rb), ("done", 1), ("act", 28)) + tuple((k, 7 * rb) for k in INFO_KEYS) + \
                    (("reward_parts", 5 * rb), ("amp_metric", rb)):
                offs[name] = (o, nbytes)
                o += (nbytes + 15) & ~15
            dev = torch.zeros(o, dtype=torch.uint8, device=b.device)
            host = torch.zeros(o, dtype=torch.uint8).pin_memory()
            info = _capi.ExoInfo()
            for k in INFO_KEYS + ("reward_parts", "amp_metric"):
                setattr(info, k, dev.data_ptr() + offs[k][0])
            hv = host.numpy()
            views = {k: hv[a:a + nb].view(np.float32 if k in ("obs", "act") else (np.uint8 if k == "done" else
                                                                                  (np.float64 if rb == 8 else np.float32)))
                     for k, (a, nb) in offs.items()}
            self._out = (dev, host, offs, info, views)
        return self._out

    def step(self, action):
        dev, host, offs, info, v = self._blob()
        b = self.batch
        v["act"][:] = np.asarray(action, dtype=np.float32).reshape(7)
        a0, an = offs["act"]
        dev[a0:a0 + an].copy_(host[a0:a0 + an], non_blocking=True)                     # the one H2D copy
        base = dev.data_ptr()
        _capi.check(b.lib.exo_step(b.h, C.c_void_p(base + a0), C.c_void_p(base + offs["obs"][0]),
                                   C.c_void_p(base + offs["reward"][0]), C.c_void_p(base + offs["done"][0]), C.byref(info),
                                   b._stream()))
        host.copy_(dev, non_blocking=True)                                              # the one D2H copy
        torch.cuda.current_stream(b.device).synchronize()
        self.counts += 1
        out = {k: v[k].astype(np.float64) for k in INFO_KEYS}
        for k, x in zip(REWARD_KEYS, v["reward_parts"]):
            out[k] = float(x)
        out["amp_metric"] = float(v["amp_metric"][0])
        self.state = v["obs"].copy()
        return self.state, float(v["reward"][0]), bool(v["done"][0]), False, out

    def return_generated_tremor_data(self):
        tt = self.batch.get_field("tremor_table")[0]          # [L, axes]
        mx = np.zeros(7)
        mx[:tt.shape[1]] = tt[:self.max_count].max(0).values.cpu().numpy()
        return self.tremor_input_sequence, mx

    def return_max_length(self):
        return self.max_count

    def return_original_joint_angles(self):
        m = self.batch.motions[int(self.file_num)]
        c = self.counts
        return [m[2, c], m[3, c], m[4, c], m[0, c], m[1, c], 0, 0]

    def render(self, mode="human"):
        raise NotImplementedError("p.GUI needs a graphics stack: unavailable on this box")

    def close(self):
        self.batch.close()
