"""The training loop of Simulation/Exoskeleton_agent_train.py:110-213 with every stage on the device (BASELINE config 5):
batched env -> policy forward (+ pink exploration noise) -> replay-buffer add -> TD7 iterations at episode end.

Structure kept from the script: one env per bundled reference motion (`num_reference_motions`, 8) - here `replicas` copies
of each, all with the same tremor sequence; the replay buffer has one row per motion (`tremor_num=i`, :139); uniform random
actions until `warmup` env-steps have been collected (:125-129, :210), afterwards select_action(use_checkpoint=False,
use_exploration=True); envs that finished idle until the whole population is done; then
maybe_train_and_checkpoint(round(mean(ep_len)), mean(score)) (:207) - i.e. about one TD7 iteration per episode time step.
The per-step suppression metrics of :150-193 come from the step kernel's counters (row N3). What is not kept: the
script's console output / plots, and its per-env Python loops.

Exploration: Agent.init_episode_noise (:203-206) draws, per episode, a [7, ep_length] pink-noise table (power-law spectrum
1/f^beta, Agent/Pink_noise.py + colorednoise.py), normalises it by its absolute maximum and scales it by
hp.exploration_noise, which decays by exploration_noise / learning_steps per select_action call (:226). pink_noise() below
synthesises the same spectrum per env with torch.fft on the device.
"""
import time

import numpy as np
import torch

from .env import BatchedExoskeletonEnv
from .policy import DevicePolicy
from .td7 import DeviceTD7, Hyperparameters

SCRIPT_DEFAULTS = dict(tremor_amplitude_range=[0.1, 1.0], first_harmonics_interval=[4.0, 6.0], second_harmonics_interval=[8.0, 10.0],
                       max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.02, dr_actuator_range=0.03,
                       matrix_noise_fraction=0.1)


def pink_noise(n, action_dim, steps, beta, generator, device):
    """[n, action_dim, steps] power-law noise, spectrum ~ 1/f^beta, unit variance before the caller's normalisation
    (colorednoise.powerlaw_psd_gaussian: Gaussian spectrum scaled by f^(-beta/2), lowest bin frozen at f = 1/steps)."""
    f = torch.fft.rfftfreq(steps, device=device)
    f[0] = 1.0 / steps
    s = f.pow(-beta / 2.0)
    w = s[1:].clone()
    if steps % 2 == 0:
        w[-1] *= 0.5
    sigma = 2.0 * torch.sqrt((w ** 2).sum()) / steps
    shape = (n, action_dim, f.numel())
    re = torch.randn(shape, device=device, generator=generator) * s      # generator None: torch's global stream
    im = torch.randn(shape, device=device, generator=generator) * s
    im[..., 0] = 0.0
    re[..., 0] *= np.sqrt(2.0)
    if steps % 2 == 0:
        im[..., -1] = 0.0
        re[..., -1] *= np.sqrt(2.0)
    return torch.fft.irfft(torch.complex(re, im), n=steps, dim=-1) / sigma


def train_device(n_steps, tremor_sequence=(0, 1, 0, 1, 0, 0, 0), replicas=1, num_reference_motions=8, warmup=25000, hp=None,
                 device="cuda:0", seed=0, updates_per_episode_step=1.0, learning_steps=None, process_group=None, precision="tf32",
                 env_kwargs=None, log=None, init_seed=None):
    """Runs the loop until n_steps env-steps (of this rank) have been collected. Returns a dict: episode returns, TD7
    iterations, env-steps, wall seconds and the learner / policy / env objects' final scalars.

    replicas: copies of each reference-motion env (the script has 1); the population is replicas x num_reference_motions.
    updates_per_episode_step: TD7 iterations per episode time step (the script: 1 - with `replicas` > 1 the ratio of
    updates to collected transitions falls accordingly, which is the knob large populations need).
    process_group: data-parallel learners, one rank per GPU (DeviceTD7.set_process_group); give every rank its own `seed`
    (envs, exploration) and the same `init_seed` (networks)."""
    hp = hp if hp is not None else Hyperparameters()
    dev = torch.device(device)
    gen = torch.Generator(device=dev).manual_seed(int(seed))
    n = replicas * num_reference_motions
    # env j replays motion j % 8 and writes replay row j % 8: consecutive adds go round the rows, as the buffer's shared write
    # pointer (TD7_buffer_multi_agent.py:59-61) expects
    motion = np.tile(np.arange(num_reference_motions, dtype=np.int32), replicas)
    kw = dict(SCRIPT_DEFAULTS)
    kw.update(env_kwargs or {})
    env = BatchedExoskeletonEnv(motion, np.asarray(tremor_sequence, np.uint8), np.arange(n, dtype=np.uint32) + 1000003 * int(seed),
                                device=dev, precision="fp32", auto_reset=False, metrics=True, **kw)
    agent = DeviceTD7(80, 7, 1, learning_steps=learning_steps or n_steps, hp=hp, env_num=num_reference_motions, device=dev, seed=seed,
                      init_seed=seed if init_seed is None else init_seed, process_group=process_group)
    policy = DevicePolicy(dev, max_action=1.0, precision=precision)
    agent.export_policy(policy)
    tremor_num = torch.from_numpy(motion).to(dev)
    ep_length = int(env.motion_lengths.max())
    exploration = float(hp.exploration_noise)
    # the reference lowers the exploration noise once per select_action call = once per LIVE env per step (select_action,
    # TD7_multi_agent_Pink_noise.py:208-228, is called per env, Exoskeleton_agent_train.py:125-129); episode t of motion m is live
    # for steps t < L_m - 3, which the host knows without a device sync
    ep_steps = (env.motion_lengths[motion].astype(np.int64) - 3)
    live_per_step = np.array([int((ep_steps > t).sum()) for t in range(int(env.motion_lengths.max()))], dtype=np.int64)
    steps_count, allow_train = 0, False
    out = dict(episode_return=[], episode_len=[], train_iterations=0, env_steps=0)
    t0 = time.perf_counter()
    while steps_count < n_steps:
        obs = env.reset()
        score = torch.zeros(n, device=dev, dtype=torch.float32)
        ep_len = torch.ones(n, device=dev, dtype=torch.float32)
        done_prev = torch.zeros(n, device=dev, dtype=torch.bool)
        noise = None
        if allow_train:
            noise = pink_noise(n, 7, ep_length, hp.beta, gen, dev)
            noise = noise / noise.abs().amax(dim=(1, 2), keepdim=True) * exploration
        for t in range(ep_length):
            if allow_train:
                act = (policy.forward(obs, use_checkpoint=False) + noise[:, :, t]).clamp_(-1.0, 1.0)
                exploration -= agent.action_noise_decrease * int(live_per_step[t])
            else:
                act = torch.rand(n, 7, device=dev, generator=gen) * 2.0 - 1.0
            nxt, rew, done = env.step(act)
            live = ~done_prev
            agent.replay_buffer.add_batch(obs, act, nxt, rew, done, tremor_num=tremor_num, skip=done_prev)
            score += rew.float() * live
            ep_len += live
            done_prev = done_prev | done
            obs = nxt
        finished = int(done_prev.sum())                    # one read-back per episode
        if finished != n:
            raise RuntimeError("%d of %d envs did not finish within the longest reference motion" % (n - finished, n))
        steps_count += int(ep_len.sum()) - n
        mean_len, mean_score = float(ep_len.mean()), float(score.mean())
        if process_group is not None:
            # data-parallel learners must take the same train / checkpoint decisions (:315-332 compares the episode return with
            # the checkpoint's): every rank uses the means over all ranks
            import torch.distributed as dist
            grp = None if process_group is True else process_group
            v = torch.tensor([mean_len, mean_score], dtype=torch.float64, device=dev)
            dist.all_reduce(v, group=grp)
            mean_len, mean_score = (float(x) / dist.get_world_size(grp) for x in v)
        out["episode_return"].append(mean_score)
        out["episode_len"].append(mean_len)
        before = agent.training_steps
        agent.maybe_train_and_checkpoint(ep_timesteps=int(round(mean_len * updates_per_episode_step)), ep_return=mean_score)
        out["train_iterations"] += agent.training_steps - before
        agent.export_policy(policy)
        if steps_count > warmup:
            allow_train = True
        if log is not None:
            log("episode %d: steps %d, mean return %.3f, TD7 iterations %d" % (len(out["episode_return"]), steps_count, mean_score,
                                                                             out["train_iterations"]))
    torch.cuda.synchronize(dev)
    out["env_steps"] = steps_count
    out["seconds"] = time.perf_counter() - t0
    out["scalars"] = agent.scalars()
    out["stats"] = env.stats() if hasattr(env, "stats") else None
    out["agent"], out["policy"], out["env"] = agent, policy, env
    return out
