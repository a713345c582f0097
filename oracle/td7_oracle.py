"""CPU oracle of the TD7 update (test infrastructure only: imported by tests/, __graft_entry__.smoke() and bench.py's
CPU legs, never by the product).

numpy restatement - forward passes, hand-derived backward passes and torch.optim.Adam - of one Agent.train() iteration
(Agent/TD7_multi_agent_Pink_noise.py):
  Encoder.zs / .zsa       :101-111
  Critic.forward          :128-144
  Actor.forward           :79-84
  AvgL1Norm, LAP_huber    :59-64
  Agent.train             :230-312   encoder update, critic update, priorities, actor update every policy_freq,
                                     target / fixed-encoder roll-over every target_update_rate
  Adam(lr, weight_decay=1e-7) :157-166 (torch defaults: betas 0.9/0.999, eps 1e-8, L2 weight decay added to the gradient)
Pinned against the reference's own Agent.train() (torch autograd, CPU) by oracle/gen_golden_td7.py ->
tests/golden/ref_td7.npz; tests/test_td7.py::test_oracle_matches_reference_train.

The batch (LAP.sample's five tensors) and the standard-normal draws behind the target-policy noise are inputs, so the
oracle, the reference and the CUDA path can be run on the same numbers. dtype float32 reproduces the reference's
arithmetic up to summation order; float64 is the tight reference for gradient tests.
"""
import copy

import numpy as np

ENC_LAYERS = ("zs1", "zs2", "zs3", "zsa1", "zsa2", "zsa3")
CRITIC_LAYERS = ("q01", "q1", "q2", "q3", "q02", "q4", "q5", "q6")
ACTOR_LAYERS = ("l0", "l1", "l2", "l3")


def layer_shapes(state_dim, action_dim, zs_dim, enc_hdim, critic_hdim, actor_hdim):
    """(out, in) of every nn.Linear, per network, in state_dict order."""
    S, A, Z, H, C, Ha = state_dim, action_dim, zs_dim, enc_hdim, critic_hdim, actor_hdim
    return {
        "encoder": [("zs1", H, S), ("zs2", H, H), ("zs3", Z, H), ("zsa1", H, Z + A), ("zsa2", H, H), ("zsa3", Z, H)],
        "critic": [("q01", C, S + A), ("q1", C, 2 * Z + C), ("q2", C, C), ("q3", 1, C),
                   ("q02", C, S + A), ("q4", C, 2 * Z + C), ("q5", C, C), ("q6", 1, C)],
        "actor": [("l0", Ha, S), ("l1", Ha, Z + Ha), ("l2", Ha, Ha), ("l3", A, Ha)],
    }


def elu(x):
    return np.where(x > 0, x, np.expm1(np.minimum(x, 0))).astype(x.dtype)


def d_elu(y):          # derivative written with the OUTPUT y = elu(x): 1 for x > 0, y + 1 otherwise
    return np.where(y > 0, 1, y + 1).astype(y.dtype)


def avgl1_fwd(x, eps=1e-8):
    d = np.maximum(np.mean(np.abs(x), axis=-1, keepdims=True), x.dtype.type(eps))
    return x / d


def avgl1_bwd(dy, x, eps=1e-8):
    n = x.shape[-1]
    m = np.mean(np.abs(x), axis=-1, keepdims=True)
    d = np.maximum(m, x.dtype.type(eps))
    live = (m >= eps).astype(x.dtype)          # clamp(min) passes the gradient where the input is >= min
    s = np.sum(dy * x, axis=-1, keepdims=True)
    return dy / d - live * np.sign(x) * s / (n * d * d)


class Net(object):
    """A bag of nn.Linear parameters {'<layer>.weight', '<layer>.bias'} with gradient slots."""

    def __init__(self, params, dtype):
        self.p = {k: np.array(v, dtype=dtype) for k, v in params.items()}
        self.g = {k: np.zeros_like(v) for k, v in self.p.items()}

    def lin(self, name, x):
        return x @ self.p[name + ".weight"].T + self.p[name + ".bias"]

    def lin_bwd(self, name, dy, x, need_dx=True, need_dw=True):
        if need_dw:
            self.g[name + ".weight"] += dy.T @ x
            self.g[name + ".bias"] += dy.sum(0)
        return dy @ self.p[name + ".weight"] if need_dx else None

    def zero_grad(self):
        for v in self.g.values():
            v[...] = 0

    def clone(self):
        return copy.deepcopy(self)


class Adam(object):
    """torch.optim.Adam, single-tensor path: L2 weight decay folded into the gradient, bias-corrected step."""

    def __init__(self, net, lr, weight_decay=1e-7, betas=(0.9, 0.999), eps=1e-8):
        self.net, self.lr, self.wd, self.betas, self.eps = net, lr, weight_decay, betas, eps
        self.m = {k: np.zeros_like(v) for k, v in net.p.items()}
        self.v = {k: np.zeros_like(v) for k, v in net.p.items()}
        self.t = 0

    def step(self):
        self.t += 1
        b1, b2 = self.betas
        bc1 = 1 - b1 ** self.t
        bc2 = 1 - b2 ** self.t
        for k, p in self.net.p.items():
            T = p.dtype.type
            g = self.net.g[k] + T(self.wd) * p
            self.m[k] = self.m[k] + (g - self.m[k]) * T(1 - b1)           # lerp_, as torch
            self.v[k] = self.v[k] * T(b2) + g * g * T(1 - b2)
            denom = np.sqrt(self.v[k]) / T(np.sqrt(bc2)) + T(self.eps)
            p -= T(self.lr / bc1) * (self.m[k] / denom)


class TD7Oracle(object):
    """Agent's learner state and train() on numpy. `params`: {'actor'|'critic'|'encoder': state dict}; the target,
    fixed and checkpoint copies start equal to them, as Agent.__init__ (:151-166) deep-copies them."""

    def __init__(self, params, dtype=np.float32, discount=0.99, target_update_rate=250, policy_freq=2,
                 target_policy_noise=0.2, noise_clip=0.5, policy_noise_decrease=0.0, alpha=0.4, min_priority=1.0,
                 actor_lr=3e-4, critic_lr=3e-4, encoder_lr=3e-4, weight_decay=1e-7, offline=False, lmbda=0.1):
        self.dtype = dtype
        self.actor = Net(params["actor"], dtype)
        self.critic = Net(params["critic"], dtype)
        self.encoder = Net(params["encoder"], dtype)
        self.actor_target = self.actor.clone()
        self.critic_target = self.critic.clone()
        self.fixed_encoder = self.encoder.clone()
        self.fixed_encoder_target = self.encoder.clone()
        self.checkpoint_actor = self.actor.clone()
        self.checkpoint_encoder = self.encoder.clone()
        self.actor_opt = Adam(self.actor, actor_lr, weight_decay)
        self.critic_opt = Adam(self.critic, critic_lr, weight_decay)
        self.encoder_opt = Adam(self.encoder, encoder_lr, weight_decay)
        self.discount, self.target_update_rate, self.policy_freq = discount, target_update_rate, policy_freq
        self.target_policy_noise, self.noise_clip, self.policy_noise_decrease = target_policy_noise, noise_clip, policy_noise_decrease
        self.alpha, self.min_priority = alpha, min_priority
        self.offline, self.lmbda = offline, lmbda
        self.training_steps = 0
        self.max, self.min, self.max_target, self.min_target = -1e8, 1e8, 0.0, 0.0
        self.last = {}

    # ---------------------------------------------------------------- forward passes (with the caches backward needs)
    @staticmethod
    def enc_zs(net, s):
        h1 = elu(net.lin("zs1", s))
        h2 = elu(net.lin("zs2", h1))
        z3 = net.lin("zs3", h2)
        return avgl1_fwd(z3), (s, h1, h2, z3)

    @staticmethod
    def enc_zsa(net, zs, a):
        c = np.concatenate([zs, a], 1)
        g1 = elu(net.lin("zsa1", c))
        g2 = elu(net.lin("zsa2", g1))
        return net.lin("zsa3", g2), (c, g1, g2)

    @staticmethod
    def critic_fwd(net, s, a, zsa, zs):
        sa = np.concatenate([s, a], 1)
        emb = np.concatenate([zsa, zs], 1)
        qs, cache = [], []
        for l0, l1, l2, l3 in (("q01", "q1", "q2", "q3"), ("q02", "q4", "q5", "q6")):
            u = net.lin(l0, sa)
            c = np.concatenate([avgl1_fwd(u), emb], 1)
            h1 = elu(net.lin(l1, c))
            h2 = elu(net.lin(l2, h1))
            qs.append(net.lin(l3, h2))
            cache.append((sa, u, c, h1, h2))
        return np.concatenate(qs, 1), cache

    @staticmethod
    def actor_fwd(net, s, zs):
        u = net.lin("l0", s)
        c = np.concatenate([avgl1_fwd(u), zs], 1)
        h1 = np.maximum(net.lin("l1", c), 0)
        h2 = np.maximum(net.lin("l2", h1), 0)
        return np.tanh(net.lin("l3", h2)), (s, u, c, h1, h2)

    # ---------------------------------------------------------------- backward passes
    @staticmethod
    def critic_bwd(net, cache, dq, need_dw, zs_dim):
        """Returns (d action [B,A], d zsa [B,Z]) when the inputs need gradients (the actor update)."""
        d_sa, d_zsa = 0, 0
        for h, (l0, l1, l2, l3) in enumerate((("q01", "q1", "q2", "q3"), ("q02", "q4", "q5", "q6"))):
            sa, u, c, h1, h2 = cache[h]
            C = u.shape[1]
            d = net.lin_bwd(l3, dq[:, h:h + 1], h2, True, need_dw) * d_elu(h2)
            d = net.lin_bwd(l2, d, h1, True, need_dw) * d_elu(h1)
            dc = net.lin_bwd(l1, d, c, True, need_dw)
            du = avgl1_bwd(dc[:, :C], u)
            dsa = net.lin_bwd(l0, du, sa, not need_dw, need_dw)
            if not need_dw:
                d_sa = d_sa + dsa
                d_zsa = d_zsa + dc[:, C:C + zs_dim]
        return d_sa, d_zsa

    # ---------------------------------------------------------------- Agent.train (:230-312)
    def train(self, state, action, next_state, reward, not_done, randn):
        T = self.dtype
        s, a, ns = (np.asarray(x, dtype=T) for x in (state, action, next_state))
        r, nd = np.asarray(reward, dtype=T).reshape(-1, 1), np.asarray(not_done, dtype=T).reshape(-1, 1)
        B, A = a.shape
        self.training_steps += 1
        # --- encoder (:238-246)
        enc = self.encoder
        next_zs, _ = self.enc_zs(enc, ns)
        zs, c_zs = self.enc_zs(enc, s)
        pred, c_zsa = self.enc_zsa(enc, zs, a)
        Z = zs.shape[1]
        self.last["encoder_loss"] = float(np.mean((pred - next_zs) ** 2))
        enc.zero_grad()
        d = (pred - next_zs) * T(2.0 / pred.size)
        c, g1, g2 = c_zsa
        d = enc.lin_bwd("zsa3", d, g2) * d_elu(g2)
        d = enc.lin_bwd("zsa2", d, g1) * d_elu(g1)
        d = enc.lin_bwd("zsa1", d, c)[:, :Z]
        s_, h1, h2, z3 = c_zs
        d = avgl1_bwd(d, z3)
        d = enc.lin_bwd("zs3", d, h2) * d_elu(h2)
        d = enc.lin_bwd("zs2", d, h1) * d_elu(h1)
        enc.lin_bwd("zs1", d, s_, need_dx=False)
        self.encoder_opt.step()
        # --- critic (:251-276)
        tzs, _ = self.enc_zs(self.fixed_encoder_target, ns)
        noise = np.clip(np.asarray(randn, dtype=T) * T(self.target_policy_noise), -T(self.noise_clip), T(self.noise_clip))
        self.target_policy_noise -= self.policy_noise_decrease
        na, _ = self.actor_fwd(self.actor_target, ns, tzs)
        na = np.clip(na + noise, -1, 1).astype(T)
        tzsa, _ = self.enc_zsa(self.fixed_encoder_target, tzs, na)
        qt, _ = self.critic_fwd(self.critic_target, ns, na, tzsa, tzs)
        qt = qt.min(1, keepdims=True)
        qt = r + nd * T(self.discount) * np.clip(qt, T(self.min_target), T(self.max_target))
        self.max = max(self.max, float(qt.max()))
        self.min = min(self.min, float(qt.min()))
        fzs, _ = self.enc_zs(self.fixed_encoder, s)
        fzsa, _ = self.enc_zsa(self.fixed_encoder, fzs, a)
        q, c_q = self.critic_fwd(self.critic, s, a, fzsa, fzs)
        diff = q - qt
        td = np.abs(diff)
        mp = T(self.min_priority)
        self.last["critic_loss"] = float(np.where(td < mp, 0.5 * td * td, mp * td).sum(1).mean())
        self.critic.zero_grad()
        dq = np.where(td < mp, td, mp) * np.sign(diff) / T(B)
        self.critic_bwd(self.critic, c_q, dq.astype(T), True, Z)
        self.critic_opt.step()
        # --- LAP priorities (:281-282)
        priority = np.maximum(td.max(1), mp) ** T(self.alpha)
        # --- actor (:287-298)
        if self.training_steps % self.policy_freq == 0:
            act, c_a = self.actor_fwd(self.actor, s, fzs)
            zsa, c_f = self.enc_zsa(self.fixed_encoder, fzs, act)
            q, c_q = self.critic_fwd(self.critic, s, act, zsa, fzs)
            self.last["actor_loss"] = float(-q.mean())
            dq = np.full_like(q, -1.0 / q.size)
            d_sa, d_zsa = self.critic_bwd(self.critic, c_q, dq, False, Z)
            fe = self.fixed_encoder
            c, g1, g2 = c_f
            d = fe.lin_bwd("zsa3", d_zsa, g2, True, False) * d_elu(g2)
            d = fe.lin_bwd("zsa2", d, g1, True, False) * d_elu(g1)
            d_act = fe.lin_bwd("zsa1", d, c, True, False)[:, Z:] + d_sa[:, -A:]
            if self.offline:                    # TD3+BC (:292-293): + lmbda * |Q|.mean().detach() * mse(actor, action)
                coef = T(self.lmbda) * np.abs(q).mean()
                self.last["actor_loss"] += float(coef * np.mean((act - a) ** 2))
                d_act = d_act + coef * T(2.0 / act.size) * (act - a)
            self.actor.zero_grad()
            s_, u, c, h1, h2 = c_a
            d = d_act * (1 - act * act)
            d = self.actor.lin_bwd("l3", d, h2) * (h2 > 0)
            d = self.actor.lin_bwd("l2", d, h1) * (h1 > 0)
            d = self.actor.lin_bwd("l1", d, c)[:, :u.shape[1]]
            d = avgl1_bwd(d, u)
            self.actor.lin_bwd("l0", d, s_, need_dx=False)
            self.actor_opt.step()
        # --- roll-over (:303-312)
        rolled = self.training_steps % self.target_update_rate == 0
        if rolled:
            self.actor_target = self.actor.clone()
            self.critic_target = self.critic.clone()
            self.fixed_encoder_target = self.fixed_encoder.clone()
            self.fixed_encoder = self.encoder.clone()
            self.max_target, self.min_target = self.max, self.min
        return priority.astype(T), rolled
