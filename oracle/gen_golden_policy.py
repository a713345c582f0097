#!/usr/bin/env python
"""Golden vectors for the policy forward pass, produced by the reference's OWN torch modules
(Agent/TD7_multi_agent_Pink_noise.py Actor / Encoder, imported unmodified from /root/reference) with the shipped
[1,1,1,1] checkpoint (Simulation/AGENT_NNS/[1,1,1,1]/[1,1,1,1]_checkpoint_{actor,encoder}), evaluated the way
select_action(use_checkpoint=True, use_exploration=False) does (:208-228), float32 on the CPU.

    python oracle/gen_golden_policy.py          # needs /root/reference; writes tests/golden/ref_policy.npz

Contents: the 14 parameter arrays the forward pass uses (float32), observations = 384 rows of the cfg-1 rollout golden
(real env observations) + 384 rows of N(0,1)*0.5 + 1 all-zero row (AvgL1Norm's clamp) , and the reference's zs [n,300]
and actions [n,7] for them.
"""
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("EXO_REFERENCE", "/root/reference")


def main():
    sys.path.insert(0, REF)
    import torch
    import torch.nn.functional as F
    from Agent.TD7_multi_agent_Pink_noise import Actor, Encoder, Hyperparameters
    hp = Hyperparameters()
    d = os.path.join(REF, "Simulation", "AGENT_NNS", "[1,1,1,1]", "[1,1,1,1]_")
    actor = Actor(80, 7, hp.zs_dim, hp.actor_hdim, hp.actor_activ)
    enc = Encoder(80, 7, hp.zs_dim, hp.enc_hdim, hp.enc_activ)
    actor.load_state_dict(torch.load(d + "checkpoint_actor", map_location="cpu", weights_only=False))
    enc.load_state_dict(torch.load(d + "checkpoint_encoder", map_location="cpu", weights_only=False))
    torch.set_num_threads(1)
    g = np.load(os.path.join(REPO, "tests", "golden", "ref_cfg1_trained_actor.npz"))
    rs = np.random.RandomState(7)
    obs = np.concatenate([g["A0/obs"][::2][:384], (rs.randn(384, 80) * 0.5).astype(np.float32), np.zeros((1, 80), np.float32)]).astype(np.float32)
    with torch.no_grad():
        s = torch.from_numpy(obs).float()
        zs = enc.zs(s)
        act = actor(s, zs).cpu().data.numpy().clip(-1, 1)
    out = {"obs": obs, "zs": zs.numpy().astype(np.float32), "action": act.astype(np.float32)}
    for k in ("zs1", "zs2", "zs3"):
        out[k + ".weight"] = enc.state_dict()[k + ".weight"].numpy()
        out[k + ".bias"] = enc.state_dict()[k + ".bias"].numpy()
    for k in ("l0", "l1", "l2", "l3"):
        out[k + ".weight"] = actor.state_dict()[k + ".weight"].numpy()
        out[k + ".bias"] = actor.state_dict()[k + ".bias"].numpy()
    path = os.path.join(REPO, "tests", "golden", "ref_policy.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", obs.shape[0], "observations; |action| max", float(np.abs(act).max()))


if __name__ == "__main__":
    main()
