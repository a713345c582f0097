"""Host-side mirror of the learner half of the reference's Agent (Agent/TD7_multi_agent_Pink_noise.py:147-366), backed
by the TD7 update kernels of libexo_env.so (include/exo_td7.h, csrc/exo_td7.cu).

  Hyperparameters   the reference's dataclass (:12-55), same fields and defaults (the activations are fixed to the
                    reference's: ELU encoder / critic, ReLU actor).
  DeviceTD7         Agent's nine networks, three Adam states and the replay buffer (exo_b200.LAP) resident on one B200;
                    train() = Agent.train (:230-312) with no host copy and no synchronisation; maybe_train_and_checkpoint
                    / train_and_reset as :315-344; save / load read and write the reference's files (:348-366:
                    <filename>_{actor,critic,encoder}[_optimizer], _checkpoint_{actor,encoder}), so a run can move
                    between the reference and this library in either direction.

PyTorch supplies device memory and streams (and pickles the state dicts); every matrix product, gradient and Adam step
runs in the CUDA kernels. There is no CPU path.
"""
import ctypes as C
from collections import OrderedDict
from dataclasses import dataclass

import numpy as np
import torch

from . import _capi
from .lap import LAP, _from_device_pointer

NETS = ("actor", "actor_target", "checkpoint_actor", "critic", "critic_target", "encoder", "fixed_encoder",
        "fixed_encoder_target", "checkpoint_encoder")
OPTS = ("actor", "critic", "encoder")


@dataclass
class Hyperparameters:
    # Generic
    batch_size: int = 128
    buffer_size: int = 2.5e5
    discount: float = 0.99
    target_update_rate: int = 250
    exploration_noise: float = 0.1
    # Pink noise
    beta: float = 1
    noise_scale: float = 0.3
    # TD3
    target_policy_noise: float = 0.2
    noise_clip: float = 0.5
    policy_freq: int = 2
    # LAP
    alpha: float = 0.4
    min_priority: float = 1
    # TD3+BC
    lmbda: float = 0.1
    # Checkpointing
    max_eps_when_checkpointing: int = 20
    steps_before_checkpointing: int = 75e4
    reset_weight: float = 0.9
    # Encoder / critic / actor models
    zs_dim: int = 300
    enc_hdim: int = 300
    encoder_lr: float = 3e-4
    critic_hdim: int = 320
    critic_lr: float = 3e-4
    actor_hdim: int = 300
    actor_lr: float = 3e-4


def _kind(net):
    return "actor" if "actor" in net else ("critic" if "critic" in net else "encoder")


def layer_shapes(state_dim, action_dim, hp):
    """(name, out, in) of every nn.Linear per network kind, in state_dict order (Actor :70-76, Critic :118-126, Encoder :92-99)."""
    S, A, Z, H, Cd, Ha = state_dim, action_dim, hp.zs_dim, hp.enc_hdim, hp.critic_hdim, hp.actor_hdim
    return {"actor": [("l0", Ha, S), ("l1", Ha, Z + Ha), ("l2", Ha, Ha), ("l3", A, Ha)],
            "critic": [("q01", Cd, S + A), ("q1", Cd, 2 * Z + Cd), ("q2", Cd, Cd), ("q3", 1, Cd),
                       ("q02", Cd, S + A), ("q4", Cd, 2 * Z + Cd), ("q5", Cd, Cd), ("q6", 1, Cd)],
            "encoder": [("zs1", H, S), ("zs2", H, H), ("zs3", Z, H), ("zsa1", H, Z + A), ("zsa2", H, H), ("zsa3", Z, H)]}


def flatten(state_dict, shapes):
    """state dict -> the flat float32 vector of include/exo_td7.h (weight [out,in] row-major, then bias, per layer)."""
    parts = []
    for name, out, inp in shapes:
        for key, shp in ((name + ".weight", (out, inp)), (name + ".bias", (out,))):
            if key not in state_dict:
                raise KeyError("state dict has no %r" % key)
            v = state_dict[key]
            v = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
            if tuple(v.shape) != shp:
                raise _capi.ExoError("%s has shape %s, this learner was built for %s" % (key, tuple(v.shape), shp))
            parts.append(np.ascontiguousarray(v, dtype=np.float32).reshape(-1))
    return np.concatenate(parts)


def unflatten(vec, shapes):
    out, o = OrderedDict(), 0
    for name, rows, cols in shapes:
        out[name + ".weight"] = torch.from_numpy(vec[o:o + rows * cols].reshape(rows, cols).copy())
        o += rows * cols
        out[name + ".bias"] = torch.from_numpy(vec[o:o + rows].copy())
        o += rows
    return out


class DeviceTD7(object):
    """Agent's learner on one B200. Constructor arguments as Agent.__init__ (:148-149) plus the device."""

    def __init__(self, state_dim, action_dim, max_action=1, learning_steps=6000000, offline=False, hp=None, env_num=15,
                 device="cuda:0", seed=0, replay_buffer=True, init_seed=None, process_group=None):
        self.offline = bool(offline)
        self.hp = hp if hp is not None else Hyperparameters()
        hp = self.hp
        self.lib = _capi.load()
        if not torch.cuda.is_available():
            raise _capi.ExoError("no CUDA device: the TD7 update runs on a B200 (there is no CPU fallback)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _capi.ExoError("device must be a CUDA device, got %s" % device)
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        self.state_dim, self.action_dim, self.max_action, self.env_num = int(state_dim), int(action_dim), max_action, int(env_num)
        self.batch_rows = self.env_num * int(hp.batch_size)
        self.policy_noise_decrease = hp.target_policy_noise / learning_steps          # :191
        self.action_noise_decrease = hp.exploration_noise / learning_steps            # :190
        self.shapes = layer_shapes(self.state_dim, self.action_dim, hp)
        cfg = _capi.ExoTD7Config(
            state_dim=self.state_dim, action_dim=self.action_dim, zs_dim=hp.zs_dim, enc_hdim=hp.enc_hdim, critic_hdim=hp.critic_hdim,
            actor_hdim=hp.actor_hdim, batch_rows=self.batch_rows, target_update_rate=int(hp.target_update_rate),
            policy_freq=int(hp.policy_freq), discount=hp.discount, target_policy_noise=hp.target_policy_noise, noise_clip=hp.noise_clip,
            policy_noise_decrease=self.policy_noise_decrease, alpha=hp.alpha, min_priority=hp.min_priority, actor_lr=hp.actor_lr,
            critic_lr=hp.critic_lr, encoder_lr=hp.encoder_lr, weight_decay=1e-7, seed=int(seed), offline=int(self.offline), lmbda=hp.lmbda)
        self.h = C.c_void_p()
        with torch.cuda.device(self.device):
            _capi.check(self.lib.exo_td7_create(idx, C.byref(cfg), C.byref(self.h)))
        self.replay_buffer = None
        if replay_buffer:
            self.replay_buffer = LAP(self.state_dim, self.action_dim, self.device, self.env_num, hp.buffer_size, hp.batch_size,
                                     max_action, normalize_actions=True, prioritized=True)
        self._priority = torch.empty(self.batch_rows, device=self.device, dtype=torch.float32)
        # checkpointing state (:171-177)
        self.eps_since_update = 0
        self.timesteps_since_update = 0
        self.max_eps_before_update = 1
        self.min_return = 1e8
        self.best_min_return = -1e8
        if init_seed is not None:
            self.init_networks(init_seed)
        self._group, self._hook, self._hook_error = None, None, None
        self._policy, self._policy_stale = None, True       # select_action's kernel-side copy of the four networks it evaluates
        self.noise = None
        self.exploration_noise = float(hp.exploration_noise)
        self.ep_length = 300                                 # Agent.__init__'s ep_length (:149): length of the pink-noise table
        if process_group is not None:
            self.set_process_group(process_group)

    # ------------------------------------------------------------------ data-parallel learners
    def set_process_group(self, group):
        """One learner per GPU (torch.distributed, NCCL): every rank samples its own batch from its own replay-buffer shard,
        the gradients are averaged over the ranks before each Adam step (exo_td7_set_grad_hook), and the value bounds are
        max / min-reduced at roll-over - N ranks x batch_rows rows behave like one Agent.train() on N x batch_rows rows.
        `group`: a torch.distributed process group, or True for the default group. All ranks must start from the same
        networks (same init_seed or the same files)."""
        import torch.distributed as dist
        self._group = None if group is True else group
        world = dist.get_world_size(self._group)

        def hook(opt, ptr, n, stream, user):
            try:
                g = _from_device_pointer(ptr, n, self.device)
                dist.all_reduce(g, op=dist.ReduceOp.SUM, group=self._group)       # ordered after the backward pass on the current stream
                g.mul_(1.0 / world)
            except BaseException as e:                                            # a ctypes callback cannot raise
                self._hook_error = e
        self._hook = _capi.EXO_TD7_GRAD_HOOK(hook)                                # keep the thunk alive
        self._dist = dist
        _capi.check(self.lib.exo_td7_set_grad_hook(self.h, C.cast(self._hook, C.c_void_p), None))

    # ------------------------------------------------------------------ networks
    def init_networks(self, seed=0):
        """Fresh networks with torch.nn.Linear's default initialisation (what Actor/Critic/Encoder.__init__ get), then the
        deep copies of Agent.__init__ (:151-163)."""
        g = torch.Generator().manual_seed(int(seed))
        for kind, net in (("actor", "actor"), ("critic", "critic"), ("encoder", "encoder")):
            sd = OrderedDict()
            for name, out, inp in self.shapes[kind]:
                bound = 1.0 / np.sqrt(inp)
                sd[name + ".weight"] = (torch.rand(out, inp, generator=g) * 2 - 1) * bound
                sd[name + ".bias"] = (torch.rand(out, generator=g) * 2 - 1) * bound
            self.load_state_dict(net, sd)
        self.sync_copies()

    def sync_copies(self):
        """actor_target, checkpoint_actor <- actor; critic_target <- critic; the three encoder copies <- encoder."""
        for dst, src in (("actor_target", "actor"), ("checkpoint_actor", "actor"), ("critic_target", "critic"),
                         ("fixed_encoder", "encoder"), ("fixed_encoder_target", "encoder"), ("checkpoint_encoder", "encoder")):
            self.copy_net(dst, src)

    def copy_net(self, dst, src):
        _capi.check(self.lib.exo_td7_copy_net(self.h, NETS.index(dst), NETS.index(src)))
        self._policy_stale = True

    def load_state_dict(self, net, state_dict):
        v = flatten(state_dict, self.shapes[_kind(net)])
        _capi.check(self.lib.exo_td7_set_net(self.h, NETS.index(net), v.ctypes.data_as(C.c_void_p), v.size))
        self._policy_stale = True

    def state_dict(self, net):
        n = int(self.lib.exo_td7_net_size(self.h, NETS.index(net)))
        v = np.empty(n, np.float32)
        _capi.check(self.lib.exo_td7_get_net(self.h, NETS.index(net), v.ctypes.data_as(C.c_void_p), n))
        return unflatten(v, self.shapes[_kind(net)])

    def _opt_vec(self, opt, which):
        n = int(self.lib.exo_td7_net_size(self.h, NETS.index(opt)))
        v, step = np.empty(n, np.float32), C.c_int64()
        _capi.check(self.lib.exo_td7_get_opt(self.h, OPTS.index(opt), which, v.ctypes.data_as(C.c_void_p), n, C.byref(step)))
        return v, int(step.value)

    def gradients(self, opt):
        """The gradient of the last backward pass of `opt`'s network, as a state-dict-shaped dict (tests)."""
        return unflatten(self._opt_vec(opt, 2)[0], self.shapes[opt])

    def optimizer_state_dict(self, opt):
        """torch.optim.Adam.state_dict() of the matching reference optimiser (:152,156,160)."""
        m, step = self._opt_vec(opt, 0)
        v, _ = self._opt_vec(opt, 1)
        md, vd = unflatten(m, self.shapes[opt]), unflatten(v, self.shapes[opt])
        keys = list(md)
        state = {} if step == 0 else {i: {"step": torch.tensor(float(step)), "exp_avg": md[k], "exp_avg_sq": vd[k]} for i, k in enumerate(keys)}
        lr = {"actor": self.hp.actor_lr, "critic": self.hp.critic_lr, "encoder": self.hp.encoder_lr}[opt]
        group = {"lr": lr, "betas": (0.9, 0.999), "eps": 1e-08, "weight_decay": 1e-07, "amsgrad": False, "maximize": False,
                 "foreach": None, "capturable": False, "differentiable": False, "fused": None, "decoupled_weight_decay": False,
                 "params": list(range(len(keys)))}
        return {"state": state, "param_groups": [group]}

    def load_optimizer_state_dict(self, opt, sd):
        keys = [n + s for n, _, _ in self.shapes[opt] for s in (".weight", ".bias")]
        st = sd.get("state", {})
        step = 0
        if st:
            step = int(float(st[0]["step"]))
            m = flatten(OrderedDict((k, st[i]["exp_avg"]) for i, k in enumerate(keys)), self.shapes[opt])
            v = flatten(OrderedDict((k, st[i]["exp_avg_sq"]) for i, k in enumerate(keys)), self.shapes[opt])
        else:
            n = int(self.lib.exo_td7_net_size(self.h, NETS.index(opt)))
            m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
        for which, vec in ((0, m), (1, v)):
            _capi.check(self.lib.exo_td7_set_opt(self.h, OPTS.index(opt), which, vec.ctypes.data_as(C.c_void_p), vec.size, step))

    # ------------------------------------------------------------------ training
    def train_on_batch(self, state, action, next_state, reward, not_done, randn=None, out=None):
        """One Agent.train() iteration (:230-312) on LAP.sample's five tensors (float32, on the device). randn
        [batch_rows, action_dim]: the standard-normal draws of :254 (None: drawn on the device). Returns (priority
        [batch_rows] on the device, rolled_over); nothing is copied to the host and the stream is not synchronised."""
        B = self.batch_rows
        ts = []
        for x, cols in ((state, self.state_dim), (action, self.action_dim), (next_state, self.state_dim), (reward, 1), (not_done, 1)):
            if not isinstance(x, torch.Tensor) or x.device != self.device or x.dtype != torch.float32 or x.numel() != B * cols:
                raise _capi.ExoError("batch tensors must be float32 on %s with %d rows" % (self.device, B))
            ts.append(x.contiguous())
        rn = None
        if randn is not None:
            rn = torch.as_tensor(randn, device=self.device, dtype=torch.float32).contiguous()
            if rn.numel() != B * self.action_dim:
                raise _capi.ExoError("randn must be [batch_rows, action_dim]")
        out = self._priority if out is None else out
        rolled = C.c_int32(0)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _capi.check(self.lib.exo_td7_train(self.h, *[C.c_void_p(t.data_ptr()) for t in ts], C.c_void_p(rn.data_ptr()) if rn is not None else None,
                                           C.c_void_p(out.data_ptr()), C.byref(rolled), stream))
        self._policy_stale = True
        if self._hook_error is not None:
            e, self._hook_error = self._hook_error, None
            raise e
        if rolled.value and self._hook is not None:
            self._reduce_value_bounds()
        return out, bool(rolled.value)

    def _reduce_value_bounds(self):
        """max / min over the ranks at roll-over (:311-312 for the union of the ranks' batches). Synchronises; once per
        target_update_rate iterations."""
        sc = self.scalars()
        v = torch.tensor([sc["max"], -sc["min"]], dtype=torch.float64, device=self.device)
        self._dist.all_reduce(v, op=self._dist.ReduceOp.MAX, group=self._group)
        mx, mn = float(v[0]), -float(v[1])
        self.set_scalars(mx, mn, mx, mn, sc["target_policy_noise"], sc["training_steps"])

    def train(self):
        """Agent.train: sample -> update -> priorities (-> reset_max_priority on roll-over)."""
        if self.replay_buffer is None:
            raise _capi.ExoError("this learner was created without a replay buffer: use train_on_batch")
        batch = self.replay_buffer.sample()
        priority, rolled = self.train_on_batch(*batch)
        self.replay_buffer.update_priority(priority)
        if rolled:
            self.replay_buffer.reset_max_priority()

    @property
    def training_steps(self):
        return int(self.scalars()["training_steps"])

    def scalars(self):
        """Synchronises. max / min / max_target / min_target (:179-182), target_policy_noise, training_steps, last losses."""
        out = (C.c_double * 9)()
        _capi.check(self.lib.exo_td7_scalars(self.h, out))
        return dict(zip(("max", "min", "max_target", "min_target", "target_policy_noise", "training_steps", "encoder_loss",
                         "critic_loss", "actor_loss"), list(out)))

    def set_scalars(self, max, min, max_target, min_target, target_policy_noise, training_steps):
        arr = (C.c_double * 6)(max, min, max_target, min_target, target_policy_noise, training_steps)
        _capi.check(self.lib.exo_td7_set_scalars(self.h, arr))

    # ------------------------------------------------------------------ checkpointing (:315-344)
    def maybe_train_and_checkpoint(self, ep_timesteps, ep_return):
        self.eps_since_update += 1
        self.timesteps_since_update += ep_timesteps
        self.min_return = min(self.min_return, ep_return)
        if self.min_return < self.best_min_return:            # end the evaluation of the current policy early
            self.train_and_reset()
        elif self.eps_since_update == self.max_eps_before_update:
            self.best_min_return = self.min_return
            self.copy_net("checkpoint_actor", "actor")
            self.copy_net("checkpoint_encoder", "fixed_encoder")
            self.train_and_reset()

    def train_and_reset(self):
        steps = self.training_steps
        for _ in range(int(self.timesteps_since_update)):
            if steps == self.hp.steps_before_checkpointing:
                self.best_min_return *= self.hp.reset_weight
                self.max_eps_before_update = self.hp.max_eps_when_checkpointing
            self.train()
            steps += 1
        self.eps_since_update = 0
        self.timesteps_since_update = 0
        self.min_return = 1e8

    def reset_buffer(self):
        self.replay_buffer.reset_buffer()

    # ------------------------------------------------------------------ files (:348-366)
    def save(self, filename):
        for net in ("critic", "actor", "encoder"):
            torch.save(self.state_dict(net), filename + "_" + net)
            torch.save(self.optimizer_state_dict(net), filename + "_" + net + "_optimizer")
        torch.save(self.state_dict("checkpoint_actor"), filename + "_checkpoint_actor")
        torch.save(self.state_dict("checkpoint_encoder"), filename + "_checkpoint_encoder")

    def load(self, filename):
        def rd(suffix):
            return torch.load(filename + suffix, map_location="cpu", weights_only=False)
        for net in ("critic", "actor", "encoder"):
            self.load_state_dict(net, rd("_" + net))
            self.load_optimizer_state_dict(net, rd("_" + net + "_optimizer"))
        self.copy_net("critic_target", "critic")
        self.copy_net("actor_target", "actor")
        self.copy_net("fixed_encoder", "encoder")
        self.copy_net("fixed_encoder_target", "encoder")
        self.load_state_dict("checkpoint_actor", rd("_checkpoint_actor"))
        self.load_state_dict("checkpoint_encoder", rd("_checkpoint_encoder"))

    # ------------------------------------------------------------------ acting (:203-228)
    def init_episode_noise(self):
        """Agent.init_episode_noise (:203-206): a [action_dim, ep_length] pink-noise table, normalised by its absolute maximum
        and scaled by the current exploration_noise."""
        from .train import pink_noise
        n = pink_noise(1, self.action_dim, self.ep_length, self.hp.beta, None, torch.device("cpu"))[0].numpy()
        self.noise = n / np.max(np.abs(n)) * self.exploration_noise

    def select_action(self, state, timestep=None, first_step=True, use_checkpoint=False, use_exploration=True):
        """Agent.select_action (:208-228), numpy in / numpy out, evaluated by the policy kernel (state 80 / hidden 300 /
        action 7 only). The kernel's copy of the networks is refreshed when an update, load or copy has changed them. For the
        batched env keep the observations on the device and call a DevicePolicy directly (export_policy)."""
        from .policy import DevicePolicy
        if self._policy is None:
            self._policy = DevicePolicy(self.device, max_action=self.max_action, precision="fp32")
        if self._policy_stale:
            self.export_policy(self._policy)
            self._policy_stale = False
        if use_exploration:
            if first_step or self.noise is None:
                self.init_episode_noise()
            self.exploration_noise -= self.action_noise_decrease
            self._policy.noise = self.noise
        return self._policy.select_action(state, timestep=timestep, first_step=first_step, use_checkpoint=use_checkpoint,
                                          use_exploration=use_exploration)

    def export_policy(self, policy):
        """Push (actor, fixed_encoder) and (checkpoint_actor, checkpoint_encoder) into a DevicePolicy - the pairs
        select_action evaluates (:210-217). Only for the 80 / 300 / 7 sizes the policy kernel is built for."""
        policy.set_networks(self.state_dict("actor"), self.state_dict("fixed_encoder"), "current")
        policy.set_networks(self.state_dict("checkpoint_actor"), self.state_dict("checkpoint_encoder"), "checkpoint")

    def device_randn(self, count, step):
        """The standard normals exo_td7_train draws for iteration `step` when it is given none (tests)."""
        out = torch.empty(int(count), device=self.device, dtype=torch.float32)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _capi.check(self.lib.exo_td7_randn(self.h, C.c_void_p(out.data_ptr()), int(count), int(step), stream))
        return out

    def launch_count(self):
        return int(self.lib.exo_td7_launch_count(self.h))

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.exo_td7_destroy(self.h)
            self.h = None
        if getattr(self, "replay_buffer", None) is not None:
            self.replay_buffer.close()
        if getattr(self, "_policy", None) is not None:
            self._policy.close()
            self._policy = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
