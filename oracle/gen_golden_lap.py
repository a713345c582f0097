#!/usr/bin/env python
"""Golden vectors for the replay buffer: runs the reference's OWN `LAP` class (Agent/TD7_buffer_multi_agent.py,
imported unmodified from /root/reference, torch on CPU) through a scripted sequence of add / sample /
update_priority / reset_max_priority calls and records every input, the torch.rand draws, the sampled indices and
outputs, the counters and the final storage.

    python oracle/gen_golden_lap.py            ->  tests/golden/ref_lap.npz

The script (phases): 40 lock-step env steps with some envs idle (the training loop's `if not done[i]`), two
sample/update rounds, reset_max_priority, 60 more steps that wrap the 64-slot buffer, one more sample/update round.
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("EXO_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
from Agent.TD7_buffer_multi_agent import LAP  # noqa: E402

E, MAX, B, SD, AD = 8, 64, 16, 80, 7


def main():
    rs = np.random.RandomState(7)
    buf = LAP(SD, AD, torch.device("cpu"), E, max_size=MAX, batch_size=B, max_action=2.0, normalize_actions=True)
    out = {"meta": np.array([E, MAX, B, SD, AD]), "max_action": np.array(2.0)}
    ev = []      # event log: (kind, index into the per-kind arrays)
    adds, samples, updates, counters = [], [], [], []

    def snap():
        counters.append([buf.ptr, buf.count, buf.size, buf.max_priority])

    def do_steps(nsteps, idle_prob):
        for _ in range(nsteps):
            st = rs.uniform(-1, 1, (E, SD)).astype(np.float32)
            ac = rs.uniform(-2, 2, (E, AD)).astype(np.float32)
            ns = rs.uniform(-1, 1, (E, SD)).astype(np.float32)
            rw = rs.uniform(0, 1, E).astype(np.float32)
            dn = rs.uniform(0, 1, E) < 0.1
            skip = rs.uniform(0, 1, E) < idle_prob
            for i in range(E):
                if not skip[i]:
                    buf.add(st[i], ac[i], ns[i], float(rw[i]), bool(dn[i]), tremor_num=i)
            adds.append(dict(state=st, action=ac, next_state=ns, reward=rw, done=dn, skip=skip))
            ev.append(0)
            snap()

    def do_sample(seed):
        torch.manual_seed(seed)
        state0 = torch.get_rng_state()
        u = torch.stack([torch.rand(size=(B,)) for _ in range(E)]).numpy()      # the draws sample() is about to make
        torch.set_rng_state(state0)
        res = buf.sample()
        ind = np.stack([np.asarray(x) for x in buf.priority_indexes])
        samples.append(dict(u=u, ind=ind, state=res[0].numpy(), action=res[1].numpy(), next_state=res[2].numpy(),
                            reward=res[3].numpy(), not_done=res[4].numpy()))
        ev.append(1)
        pr = torch.tensor(rs.uniform(0.05, 3.0, E * B).astype(np.float32))
        buf.update_priority(pr)
        updates.append(pr.numpy())
        ev.append(2)
        snap()

    do_steps(40, 0.15)
    do_sample(11)
    do_sample(12)
    buf.reset_max_priority()
    ev.append(3)
    snap()
    do_steps(60, 0.3)
    do_sample(13)

    out["events"] = np.array(ev)
    for k in adds[0]:
        out["add_" + k] = np.stack([a[k] for a in adds])
    for k in samples[0]:
        out["sample_" + k] = np.stack([s[k] for s in samples])
    out["update_priority"] = np.stack(updates)
    out["counters"] = np.array(counters, dtype=np.float64)
    out["final_state"] = buf.state.astype(np.float32)
    out["final_action"] = buf.action.astype(np.float32)
    out["final_next_state"] = buf.next_state.astype(np.float32)
    out["final_reward"] = buf.reward.astype(np.float32)
    out["final_not_done"] = buf.not_done.astype(np.float32)
    out["final_priority"] = buf.priority.numpy()
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "ref_lap.npz")
    np.savez_compressed(dst, **out)
    print("wrote", os.path.normpath(dst), "events", len(ev), "final counters", counters[-1])


if __name__ == "__main__":
    main()
