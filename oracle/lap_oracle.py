"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's replay buffer `LAP`
(Agent/TD7_buffer_multi_agent.py:5-150), the checker for csrc/exo_lap.cu. Imported by tests/ only; the product
never imports it.

Pinned against the reference's own class (imported unmodified from /root/reference by oracle/gen_golden_lap.py,
torch on CPU): tests/golden/ref_lap.npz, checked by tests/test_lap.py::test_oracle_matches_reference_golden.

Differences from the reference that do not change results: the uniforms of sample() are passed in (the reference
draws torch.rand(batch_size) per env; the golden records those draws), and storage is float32 where the reference
keeps float64 copies of float32 inputs and casts back to float32 in sample() (:105-111).
"""
import numpy as np


class LapOracle(object):
    def __init__(self, state_dim, action_dim, num_envs, max_size, batch_size, max_action=1, normalize_actions=True):
        # LAP.__init__ (:6-47)
        self.max_size, self.batch_size, self.num_envs = int(max_size), int(batch_size), int(num_envs)
        self.state_dim, self.action_dim = state_dim, action_dim
        self.ptr = self.count = self.size = 0
        self.state = np.zeros((num_envs, self.max_size, state_dim), np.float32)
        self.action = np.zeros((num_envs, self.max_size, action_dim), np.float32)
        self.next_state = np.zeros((num_envs, self.max_size, state_dim), np.float32)
        self.reward = np.zeros((num_envs, self.max_size, 1), np.float32)
        self.not_done = np.zeros((num_envs, self.max_size, 1), np.float32)
        self.priority = np.zeros((num_envs, self.max_size), np.float32)
        self.max_priority = 1.0
        self.priority_indexes = []
        self.normalize_actions = max_action if normalize_actions else 1

    def add(self, state, action, next_state, reward, done, tremor_num):
        # LAP.add (:49-63): the pointer advances only when the running count is a multiple of num_envs
        self.state[tremor_num, self.ptr] = state
        self.action[tremor_num, self.ptr] = np.asarray(action, np.float32) / np.float32(self.normalize_actions)
        self.next_state[tremor_num, self.ptr] = next_state
        self.reward[tremor_num, self.ptr] = reward
        self.not_done[tremor_num, self.ptr] = 1.0 - float(done)
        self.priority[tremor_num, self.ptr] = self.max_priority
        if self.count % self.num_envs == 0:
            self.ptr = (self.ptr + 1) % self.max_size
            self.size = min(self.size + 1, self.max_size)
        self.count += 1

    def add_batch(self, state, action, next_state, reward, done, tremor_num=None, skip=None):
        for j in range(len(state)):
            if skip is not None and skip[j]:
                continue
            self.add(state[j], action[j], next_state[j], reward[j], done[j], j % self.num_envs if tremor_num is None else int(tremor_num[j]))

    def sample(self, uniforms, exact=False):
        # LAP.sample (:65-111), prioritized branch; uniforms [num_envs, batch_size] float32 = the torch.rand draws.
        # exact=True accumulates the prefix sum in float64 (the distribution the reference's float32 torch.cumsum
        # approximates: over ~4e4 slots its running sum is off by ~1e-5 of the total, i.e. by whole slots, and the
        # error depends on the summation order, so torch-CPU, torch-CUDA and numpy already disagree with each other).
        B = self.batch_size
        out = [np.zeros((self.num_envs * B, d), np.float32) for d in (self.state_dim, self.action_dim, self.state_dim, 1, 1)]
        self.priority_indexes = []
        u = np.asarray(uniforms, np.float32).reshape(self.num_envs, B)
        for i in range(self.num_envs):
            if exact:
                csum = np.cumsum(self.priority[i, :self.size].astype(np.float64))
                val = (u[i] * np.float32(csum[-1])).astype(np.float64)
            else:
                csum = np.cumsum(self.priority[i, :self.size], dtype=np.float32)
                val = u[i] * csum[-1]
            ind = np.minimum(np.searchsorted(csum, val, side="left"), self.size - 1)
            self.priority_indexes.append(ind)
            for o, src in zip(out, (self.state, self.action, self.next_state, self.reward, self.not_done)):
                o[i * B:(i + 1) * B] = src[i, ind]
        return out

    def boundary_distance(self, uniforms):
        """|val - nearest cumsum entry| / total per sample: samples closer than float32 rounding to a boundary may
        legitimately land on the neighbouring index when the prefix sum is accumulated in another order."""
        B = self.batch_size
        u = np.asarray(uniforms, np.float32).reshape(self.num_envs, B)
        d = np.zeros((self.num_envs, B))
        for i in range(self.num_envs):
            csum = np.cumsum(self.priority[i, :self.size].astype(np.float64))
            val = (u[i] * np.float32(csum[-1])).astype(np.float64)
            d[i] = np.min(np.abs(val[:, None] - csum[None, :]), axis=1) / csum[-1]
        return d

    def update_priority(self, priority):
        # LAP.update_priority (:113-117)
        p = np.asarray(priority, np.float32).reshape(-1)
        B = self.batch_size
        for i in range(self.num_envs):
            self.priority[i, self.priority_indexes[i]] = p[i * B:(i + 1) * B]
        self.max_priority = max(float(p.max()), self.max_priority)
        self.priority_indexes = []

    def reset_max_priority(self):
        self.max_priority = float(self.priority.max())      # (:119-120)

    def reset_buffer(self):
        na = self.normalize_actions                           # (:122-139)
        self.__init__(self.state_dim, self.action_dim, self.num_envs, self.max_size, self.batch_size, na, True)
