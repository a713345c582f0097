#!/usr/bin/env python
"""Timing of the device TD7 update (exo_b200.DeviceTD7.train_on_batch) beside the same update written with torch modules,
autograd and torch.optim.Adam on the same GPU (float32, allow_tf32 off - the reference's arithmetic; this is how the
reference's Agent.train runs when it has a GPU). Prints ms per iteration for both and the parameter drift between them.

    python tools/td7_probe.py [--envs 15] [--iters 200]
"""
import argparse
import copy
import os
import sys
import time

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import exo_b200  # noqa: E402


def avg_l1(x, eps=1e-8):
    return x / x.abs().mean(-1, keepdim=True).clamp(min=eps)


class TorchNet(nn.Module):
    def __init__(self, shapes):
        super().__init__()
        for name, out, inp in shapes:
            setattr(self, name, nn.Linear(inp, out))


class TorchTD7(object):
    """The update of Agent.train written against plain torch (a timing/parity stand-in, not the reference's file)."""

    def __init__(self, learner, dev):
        sh = learner.shapes
        self.nets = {}
        for n in exo_b200.td7.NETS:
            net = TorchNet(sh[exo_b200.td7._kind(n)]).to(dev)
            net.load_state_dict(learner.state_dict(n))
            self.nets[n] = net
        hp = learner.hp
        self.opt = {k: torch.optim.Adam(self.nets[k].parameters(), lr=3e-4, weight_decay=1e-7) for k in ("actor", "critic", "encoder")}
        self.hp, self.steps = hp, 0
        self.max, self.min, self.max_target, self.min_target = -1e8, 1e8, 0.0, 0.0
        self.sigma, self.dec = hp.target_policy_noise, learner.policy_noise_decrease

    @staticmethod
    def zs(e, s):
        return avg_l1(e.zs3(F.elu(e.zs2(F.elu(e.zs1(s))))))

    @staticmethod
    def zsa(e, zs, a):
        return e.zsa3(F.elu(e.zsa2(F.elu(e.zsa1(torch.cat([zs, a], 1))))))

    @staticmethod
    def actor(n, s, zs):
        a = torch.cat([avg_l1(n.l0(s)), zs], 1)
        return torch.tanh(n.l3(F.relu(n.l2(F.relu(n.l1(a))))))

    @staticmethod
    def critic(n, s, a, zsa, zs):
        sa, emb = torch.cat([s, a], 1), torch.cat([zsa, zs], 1)
        q1 = n.q3(F.elu(n.q2(F.elu(n.q1(torch.cat([avg_l1(n.q01(sa)), emb], 1))))))
        q2 = n.q6(F.elu(n.q5(F.elu(n.q4(torch.cat([avg_l1(n.q02(sa)), emb], 1))))))
        return torch.cat([q1, q2], 1)

    def train(self, s, a, ns, r, nd, randn, sync_scalars=True):
        N, hp = self.nets, self.hp
        self.steps += 1
        with torch.no_grad():
            next_zs = self.zs(N["encoder"], ns)
        loss = F.mse_loss(self.zsa(N["encoder"], self.zs(N["encoder"], s), a), next_zs)
        self.opt["encoder"].zero_grad()
        loss.backward()
        self.opt["encoder"].step()
        with torch.no_grad():
            tzs = self.zs(N["fixed_encoder_target"], ns)
            noise = (randn * self.sigma).clamp(-hp.noise_clip, hp.noise_clip)
            self.sigma -= self.dec
            na = (self.actor(N["actor_target"], ns, tzs) + noise).clamp(-1, 1)
            tzsa = self.zsa(N["fixed_encoder_target"], tzs, na)
            qt = self.critic(N["critic_target"], ns, na, tzsa, tzs).min(1, keepdim=True)[0]
            qt = r + nd * hp.discount * qt.clamp(self.min_target, self.max_target)
            if sync_scalars:                                  # the reference reads them back every iteration (:263-264)
                self.max = max(self.max, float(qt.max()))
                self.min = min(self.min, float(qt.min()))
            fzs = self.zs(N["fixed_encoder"], s)
            fzsa = self.zsa(N["fixed_encoder"], fzs, a)
        q = self.critic(N["critic"], s, a, fzsa, fzs)
        td = (q - qt).abs()
        closs = torch.where(td < hp.min_priority, 0.5 * td.pow(2), hp.min_priority * td).sum(1).mean()
        self.opt["critic"].zero_grad()
        closs.backward()
        self.opt["critic"].step()
        prio = td.max(1)[0].clamp(min=hp.min_priority).pow(hp.alpha).detach()
        if self.steps % hp.policy_freq == 0:
            act = self.actor(N["actor"], s, fzs)
            q = self.critic(N["critic"], s, act, self.zsa(N["fixed_encoder"], fzs, act), fzs)
            aloss = -q.mean()
            self.opt["actor"].zero_grad()
            aloss.backward()
            self.opt["actor"].step()
        if self.steps % hp.target_update_rate == 0:
            N["actor_target"].load_state_dict(N["actor"].state_dict())
            N["critic_target"].load_state_dict(N["critic"].state_dict())
            N["fixed_encoder_target"].load_state_dict(N["fixed_encoder"].state_dict())
            N["fixed_encoder"].load_state_dict(N["encoder"].state_dict())
            self.max_target, self.min_target = self.max, self.min
        return prio


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=15)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--device-only", action="store_true", help="skip the torch stand-in (for ncu captures)")
    args = ap.parse_args()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    dev = torch.device("cuda:0")
    hp = exo_b200.Hyperparameters(batch_size=args.batch)
    t = exo_b200.DeviceTD7(80, 7, 1, hp=hp, env_num=args.envs, device=dev, replay_buffer=False, init_seed=1)
    ref = None if args.device_only else TorchTD7(t, dev)
    B = args.envs * args.batch
    g = torch.Generator(device=dev).manual_seed(0)
    def batch():
        return [torch.randn(B, 80, device=dev, generator=g) * 0.6, (torch.randn(B, 7, device=dev, generator=g) * 0.5).clamp(-1, 1),
                torch.randn(B, 80, device=dev, generator=g) * 0.6, torch.randn(B, 1, device=dev, generator=g) * 3 + 1,
                (torch.rand(B, 1, device=dev, generator=g) > 0.1).float(), torch.randn(B, 7, device=dev, generator=g)]
    batches = [batch() for _ in range(8)]
    # parity over a few iterations
    for i in range(6):
        p1, _ = t.train_on_batch(*batches[i])
        if ref is None:
            continue
        p2 = ref.train(*batches[i])
        print("iter %d: priority max |diff| %.2e" % (i + 1, float((p1 - p2).abs().max())))
    for n in (("actor", "critic", "encoder") if ref is not None else ()):
        sd, rd = t.state_dict(n), ref.nets[n].state_dict()
        d = torch.cat([(sd[k] - rd[k].cpu()).abs().reshape(-1) for k in sd])
        print("%-8s params after 6 iterations: max |diff| %.2e, share within 2e-6: %.5f" % (n, float(d.max()), float((d < 2e-6).float().mean())))
    out = {}
    for name, fn in (("device TD7 (libexo_env)", lambda b: t.train_on_batch(*b)), ("torch autograd + Adam", lambda b: ref.train(*b)),
                     ("torch, no per-iteration readback", lambda b: ref.train(*b, sync_scalars=False)))[:1 if ref is None else 3]:
        for i in range(10):
            fn(batches[i % 8])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(args.iters):
            fn(batches[i % 8])
        b.record()
        torch.cuda.synchronize()
        out[name] = a.elapsed_time(b) / args.iters
        print("%-36s %.3f ms per iteration (events), %.3f ms wall; %d rows" % (name, out[name], 1e3 * (time.perf_counter() - t0) / args.iters, B))
    print("launches per device iteration: %.1f" % (t.launch_count() / (6 + 10 + args.iters)))


if __name__ == "__main__":
    main()
