#!/usr/bin/env python
"""Golden vectors for the TD7 update, produced by the reference's OWN Agent.train() (Agent/TD7_multi_agent_Pink_noise.py
:230-312, torch autograd + torch.optim.Adam, float32 on the CPU, imported unmodified from /root/reference).

    python oracle/gen_golden_td7.py          # needs /root/reference; writes tests/golden/ref_td7.npz

The Agent is built with small layer sizes (Hyperparameters are the reference's own knobs) so the fixture stays small;
only the instance's replay_buffer.sample / update_priority are replaced by closures that feed a fixed batch and record
the priorities, and torch is re-seeded before every train() so the standard-normal draws of torch.randn_like(action)
(:254) can be stored beside the batch. 6 iterations with target_update_rate = 4 and policy_freq = 2: three actor
updates, one roll-over of the target / fixed networks, a decaying target_policy_noise.

Contents: dims, hyper-parameters, init/<net>/<param>, step<k>/{state,action,next_state,reward,not_done,randn,priority,
scalars = [max, min, max_target, min_target, target_policy_noise]}, after3/<net>/<param> (actor, critic, encoder) and
after6/<net>/<param> for all nine networks; off/... = the same batches through an Agent(offline=True) (TD3+BC actor loss), 4
iterations: off/init/<net>/<param>, off/step<k>/priority, off/after4/<net>/<param>.
"""
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("EXO_REFERENCE", "/root/reference")
NETS = ("actor", "actor_target", "checkpoint_actor", "critic", "critic_target", "encoder", "fixed_encoder",
        "fixed_encoder_target", "checkpoint_encoder")


def main():
    sys.path.insert(0, REF)
    import torch
    from Agent.TD7_multi_agent_Pink_noise import Agent, Hyperparameters
    torch.set_num_threads(1)
    torch.manual_seed(11)
    S, A, envs, bs, steps = 11, 3, 3, 8, 6
    hp = Hyperparameters(batch_size=bs, buffer_size=64, target_update_rate=4, zs_dim=24, enc_hdim=20, critic_hdim=28,
                         actor_hdim=16)
    agent = Agent(S, A, 1, learning_steps=100, hp=hp, env_num=envs, ep_length=16)
    B = envs * bs
    out = {"dims": np.array([S, A, hp.zs_dim, hp.enc_hdim, hp.critic_hdim, hp.actor_hdim, B], np.int64),
           "hyper": np.array([hp.discount, hp.target_update_rate, hp.policy_freq, hp.target_policy_noise, hp.noise_clip,
                              agent.policy_noise_decrease, hp.alpha, hp.min_priority, hp.actor_lr, hp.critic_lr,
                              hp.encoder_lr, 1e-7], np.float64)}
    for n in ("actor", "critic", "encoder"):
        for k, v in getattr(agent, n).state_dict().items():
            out["init/%s/%s" % (n, k)] = v.numpy().copy()
    rs = np.random.RandomState(5)
    rec = {}
    agent.replay_buffer.update_priority = lambda p: rec.__setitem__("priority", p.detach().numpy().copy())
    for t in range(1, steps + 1):
        batch = dict(state=rs.randn(B, S) * 0.7, action=np.clip(rs.randn(B, A) * 0.6, -1, 1), next_state=rs.randn(B, S) * 0.7,
                     reward=rs.randn(B, 1) * 2.0 + 0.5, not_done=(rs.rand(B, 1) > 0.15).astype(np.float64))
        batch = {k: v.astype(np.float32) for k, v in batch.items()}
        agent.replay_buffer.sample = lambda b=batch: tuple(torch.tensor(b[k], dtype=torch.float) for k in
                                                           ("state", "action", "next_state", "reward", "not_done"))
        torch.manual_seed(1000 + t)
        randn = torch.randn(B, A).numpy().copy()
        torch.manual_seed(1000 + t)
        agent.train()
        for k, v in batch.items():
            out["step%d/%s" % (t, k)] = v
        out["step%d/randn" % t] = randn
        out["step%d/priority" % t] = rec["priority"]
        out["step%d/scalars" % t] = np.array([agent.max, agent.min, agent.max_target, agent.min_target,
                                              hp.target_policy_noise], np.float64)
        if t in (3, 6):
            for n in (NETS if t == 6 else ("actor", "critic", "encoder")):
                for k, v in getattr(agent, n).state_dict().items():
                    out["after%d/%s/%s" % (t, n, k)] = v.numpy().copy()
    # the offline (TD3+BC) actor loss (:292-293): a second Agent with offline=True, 4 iterations from its own initial networks
    torch.manual_seed(12)
    hp2 = Hyperparameters(batch_size=bs, buffer_size=64, target_update_rate=4, zs_dim=24, enc_hdim=20, critic_hdim=28, actor_hdim=16)
    off = Agent(S, A, 1, learning_steps=100, offline=True, hp=hp2, env_num=envs, ep_length=16)
    for n in ("actor", "critic", "encoder"):
        for k, v in getattr(off, n).state_dict().items():
            out["off/init/%s/%s" % (n, k)] = v.numpy().copy()
    off.replay_buffer.update_priority = lambda p: rec.__setitem__("priority", p.detach().numpy().copy())
    for t in range(1, 5):
        batch = {k: out["step%d/%s" % (t, k)] for k in ("state", "action", "next_state", "reward", "not_done")}
        off.replay_buffer.sample = lambda b=batch: tuple(torch.tensor(b[k], dtype=torch.float) for k in
                                                         ("state", "action", "next_state", "reward", "not_done"))
        torch.manual_seed(1000 + t)                      # the same randn draws as step<t>/randn
        off.train()
        out["off/step%d/priority" % t] = rec["priority"]
    for n in ("actor", "critic", "encoder"):
        for k, v in getattr(off, n).state_dict().items():
            out["off/after4/%s/%s" % (n, k)] = v.numpy().copy()
    path = os.path.join(REPO, "tests", "golden", "ref_td7.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes; training_steps", agent.training_steps, "max/min", agent.max, agent.min)


if __name__ == "__main__":
    main()
