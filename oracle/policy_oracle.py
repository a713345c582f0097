"""CPU oracle of the policy forward pass (test infrastructure only: imported by tests/, __graft_entry__.smoke() and
bench.py's CPU legs, never by the product).

numpy restatement of the networks select_action evaluates (Agent/TD7_multi_agent_Pink_noise.py):
  AvgL1Norm      :59-60    x / clamp(mean(|x|, -1), min=1e-8)
  Encoder.zs     :101-105  AvgL1Norm(zs3(elu(zs2(elu(zs1(state))))))
  Actor.forward  :79-84    tanh(l3(relu(l2(relu(l1(cat[AvgL1Norm(l0(state)), zs]))))))
  select_action  :208-228  .clip(-1, 1) * max_action (exploration off)
Pinned against the reference's own torch modules with the shipped [1,1,1,1] checkpoint by oracle/gen_golden_policy.py
(tests/golden/ref_policy.npz; tests/test_policy.py::test_oracle_matches_reference_modules).

mode "fp32": float32 throughout, as torch's default matmul.
mode "tf32": the arithmetic of csrc/exo_policy.cu - weights, biases, observations and every stored activation rounded to TF32
             (round to nearest, ties away: cvt.rna.tf32.f32), products accumulated in float32 (float64 here, which
             differs from any float32 summation order by < 1e-6), bias/activation in float32.
"""
import numpy as np

KEYS = ("zs1", "zs2", "zs3", "l0", "l1", "l2", "l3")


def round_tf32(x):
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    finite = (u & np.uint32(0x7F800000)) != np.uint32(0x7F800000)
    r = ((u.astype(np.uint64) + 0x1000) & 0xFFFFE000).astype(np.uint32)
    return np.where(finite, r, u).view(np.float32)


def avg_l1_norm(x, eps=1e-8):
    return x / np.maximum(np.mean(np.abs(x), axis=-1, keepdims=True, dtype=np.float32), np.float32(eps))


def elu(x):
    return np.where(x > 0, x, np.expm1(np.minimum(x, 0))).astype(np.float32)


class PolicyOracle(object):
    def __init__(self, weights, mode="fp32"):
        """weights: dict '<layer>.weight' [out,in] / '<layer>.bias' for zs1..zs3 (encoder) and l0..l3 (actor)."""
        assert mode in ("fp32", "tf32")
        self.mode = mode
        self.w = {k: np.asarray(weights[k + ".weight"], dtype=np.float32) for k in KEYS}
        self.b = {k: np.asarray(weights[k + ".bias"], dtype=np.float32) for k in KEYS}
        if mode == "tf32":
            self.w = {k: round_tf32(v) for k, v in self.w.items()}
            self.b = {k: round_tf32(v) for k, v in self.b.items()}      # the biases ride in the GEMMs as a K column

    def _lin(self, name, x):
        if self.mode == "tf32":
            return (round_tf32(x).astype(np.float64) @ self.w[name].astype(np.float64).T).astype(np.float32) + self.b[name]
        return x @ self.w[name].T + self.b[name]

    def zs(self, state):
        state = np.asarray(state, dtype=np.float32)
        h = elu(self._lin("zs1", state))
        h = elu(self._lin("zs2", h))
        return avg_l1_norm(self._lin("zs3", h))

    def actor(self, state, zs):
        state = np.asarray(state, dtype=np.float32)
        a = avg_l1_norm(self._lin("l0", state))
        a = np.concatenate([a, zs], axis=1)
        a = np.maximum(self._lin("l1", a), 0)
        a = np.maximum(self._lin("l2", a), 0)
        return np.tanh(self._lin("l3", a)).astype(np.float32)

    def select_action(self, state, max_action=1.0):
        state = np.atleast_2d(np.asarray(state, dtype=np.float32))
        return self.actor(state, self.zs(state)).clip(-1, 1) * np.float32(max_action)
