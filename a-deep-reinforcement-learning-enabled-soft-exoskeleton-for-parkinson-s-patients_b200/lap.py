"""Host-side mirror of the reference's replay buffer, backed by the device-resident LAP of libexo_env.so.

  LAP   same constructor, attributes and methods as Agent/TD7_buffer_multi_agent.py:5-150 (add / sample /
        update_priority / reset_max_priority / reset_buffer; ptr, count, size, max_priority), so
        TD7_multi_agent*.Agent can hold it in place of its own buffer, plus add_batch() which takes the batched env's
        device tensors directly (no host copy, no synchronisation).

PyTorch supplies device memory, streams and the uniform draws of sample() (torch.rand, as the reference); storage,
pointer arithmetic, the prefix sums, the search and the gathers run in the CUDA kernels of csrc/exo_lap.cu.
"""
import ctypes as C

import numpy as np
import torch

from . import _capi


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class _DevicePointer(object):
    """float32 device memory owned by the library, exposed through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


def _from_device_pointer(ptr, n, device):
    with torch.cuda.device(device):
        return torch.as_tensor(_DevicePointer(ptr, n), device=device)


class LAP(object):
    """Prioritized (loss-adjusted) replay buffer with one row per env, resident on one B200."""

    def __init__(self, state_dim, action_dim, device, num_envs, max_size=1e6, batch_size=64, max_action=1,
                 normalize_actions=True, prioritized=True):
        self.lib = _capi.load()
        if not prioritized:
            raise NotImplementedError("only the prioritized branch of LAP.sample is built (TD7 always uses it, "
                                      "TD7_multi_agent.py:175)")
        if not torch.cuda.is_available():
            raise _capi.ExoError("no CUDA device: the replay buffer lives in B200 HBM (there is no CPU fallback)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _capi.ExoError("device must be a CUDA device, got %s" % device)
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        self.max_size = int(max_size)
        self.max_action = max_action
        self.batch_size = int(batch_size)
        self.num_envs = int(num_envs)
        self.state_dim, self.action_dim = int(state_dim), int(action_dim)
        self.prioritized = True
        self.normalize_actions = max_action if normalize_actions else 1
        self.h = C.c_void_p()
        with torch.cuda.device(idx):
            _capi.check(self.lib.exo_lap_create(self.state_dim, self.action_dim, self.num_envs, self.max_size, self.batch_size,
                                                float(max_action), int(bool(normalize_actions)), idx, C.byref(self.h)))
        n = self.num_envs * self.batch_size
        dev = self.device
        self._out = (torch.empty(n, self.state_dim, device=dev), torch.empty(n, self.action_dim, device=dev),
                     torch.empty(n, self.state_dim, device=dev), torch.empty(n, 1, device=dev), torch.empty(n, 1, device=dev))
        self._ind = torch.empty(n, dtype=torch.int32, device=dev)
        self.single_draw = False      # True: one torch.rand(num_envs * batch_size) instead of one draw per env

    # ------------------------------------------------------------------ plumbing
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.exo_lap_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _counters(self):
        p, c, s, m = C.c_int64(), C.c_int64(), C.c_int64(), C.c_double()
        _capi.check(self.lib.exo_lap_counters(self.h, C.byref(p), C.byref(c), C.byref(s), C.byref(m), self._stream()))
        return p.value, c.value, s.value, m.value

    # reference attributes (device-resident; reading one synchronises the stream)
    ptr = property(lambda self: self._counters()[0])
    count = property(lambda self: self._counters()[1])
    size = property(lambda self: self._counters()[2])
    max_priority = property(lambda self: self._counters()[3])

    @property
    def launch_count(self):
        return int(self.lib.exo_lap_launch_count(self.h))

    def storage(self, field):
        """Device view of one storage array: 'state' | 'action' | 'next_state' | 'reward' | 'not_done' | 'priority',
        shaped [num_envs, max_size, k] like the reference's arrays (priority: [num_envs, max_size])."""
        fid = ("state", "action", "next_state", "reward", "not_done", "priority").index(field)
        p, k = C.c_void_p(), C.c_int64()
        _capi.check(self.lib.exo_lap_storage(self.h, fid, C.byref(p), C.byref(k)))
        n = self.num_envs * self.max_size * k.value
        t = _from_device_pointer(p.value, n, self.device)
        return t.view(self.num_envs, self.max_size) if field == "priority" else t.view(self.num_envs, self.max_size, k.value)

    # ------------------------------------------------------------------ reference methods
    def add(self, state, action, next_state, reward, done, tremor_num):
        """LAP.add (:49-63) for one transition given as host values (numpy / floats), as the reference's training
        loop calls it. For the batched env use add_batch: this path costs one small H2D copy per call."""
        dev = self.device
        s = torch.as_tensor(np.asarray(state, dtype=np.float32).reshape(1, self.state_dim)).to(dev)
        a = torch.as_tensor(np.asarray(action, dtype=np.float32).reshape(1, self.action_dim)).to(dev)
        ns = torch.as_tensor(np.asarray(next_state, dtype=np.float32).reshape(1, self.state_dim)).to(dev)
        r = torch.tensor([float(reward)], dtype=torch.float32, device=dev)
        d = torch.tensor([1 if done else 0], dtype=torch.uint8, device=dev)
        t = torch.tensor([int(tremor_num)], dtype=torch.int32, device=dev)
        self.add_batch(s, a, ns, r, d, tremor_num=t)

    def add_batch(self, state, action, next_state, reward, done, tremor_num=None, skip=None):
        """k calls of LAP.add in index order, entirely on the device: state/next_state [k,sd], action [k,ad],
        reward [k], done [k] (bool/uint8), tremor_num [k] int32 (None: j % num_envs), skip [k] (non-zero: entry not
        added — the training loop's `if not done[i]` with the PREVIOUS step's done flags). Takes the batched env's
        output tensors as they are; nothing is copied to the host and the stream is not synchronised."""
        dev = self.device
        s = torch.as_tensor(state, device=dev, dtype=torch.float32).contiguous()
        k = s.shape[0]
        a = torch.as_tensor(action, device=dev, dtype=torch.float32).contiguous()
        ns = torch.as_tensor(next_state, device=dev, dtype=torch.float32).contiguous()
        r = torch.as_tensor(reward, device=dev).to(torch.float32).contiguous().reshape(-1)
        d = torch.as_tensor(done, device=dev)
        d = d.view(torch.uint8) if d.dtype == torch.bool else d.to(torch.uint8)
        d = d.contiguous().reshape(-1)
        if s.shape != (k, self.state_dim) or ns.shape != (k, self.state_dim) or a.shape != (k, self.action_dim) or \
                r.numel() != k or d.numel() != k:
            raise ValueError("add_batch: inconsistent shapes")
        t = None if tremor_num is None else torch.as_tensor(tremor_num, device=dev).to(torch.int32).contiguous().reshape(-1)
        sk = None
        if skip is not None:
            sk = torch.as_tensor(skip, device=dev)
            sk = (sk.view(torch.uint8) if sk.dtype == torch.bool else sk.to(torch.uint8)).contiguous().reshape(-1)
        _capi.check(self.lib.exo_lap_add(self.h, _ptr(s), _ptr(a), _ptr(ns), _ptr(r), _ptr(d), _ptr(t), _ptr(sk), k, self._stream()))

    def sample(self, uniforms=None):
        """LAP.sample (:65-111): five float32 tensors [num_envs*batch_size, dim], rows grouped by env. The uniforms are
        drawn like the reference does (torch.rand(batch_size) on the buffer's device, once per env, in env order) unless
        given; the returned tensors are reused by the next call."""
        if uniforms is None:
            if self.single_draw:
                u = torch.rand(self.num_envs * self.batch_size, device=self.device)
            else:
                u = torch.cat([torch.rand(size=(self.batch_size,), device=self.device) for _ in range(self.num_envs)])
        else:
            u = torch.as_tensor(uniforms, device=self.device, dtype=torch.float32).contiguous().reshape(-1)
            if u.numel() != self.num_envs * self.batch_size:
                raise ValueError("uniforms must have num_envs * batch_size entries")
        st, ac, ns, rw, nd = self._out
        _capi.check(self.lib.exo_lap_sample(self.h, _ptr(u), _ptr(st), _ptr(ac), _ptr(ns), _ptr(rw), _ptr(nd), _ptr(self._ind),
                                            self._stream()))
        self.ind = self._ind.view(self.num_envs, self.batch_size)      # reference: self.ind / priority_indexes
        return st, ac, ns, rw, nd

    def update_priority(self, priority):
        """LAP.update_priority (:113-117): priority [num_envs*batch_size] for the rows of the last sample()."""
        p = torch.as_tensor(priority, device=self.device).detach().to(torch.float32).contiguous().reshape(-1)
        if p.numel() != self.num_envs * self.batch_size:
            raise ValueError("priority must have num_envs * batch_size entries")
        _capi.check(self.lib.exo_lap_update_priority(self.h, _ptr(p), self._stream()))

    def reset_max_priority(self):
        _capi.check(self.lib.exo_lap_reset_max_priority(self.h, self._stream()))

    def reset_buffer(self):
        _capi.check(self.lib.exo_lap_reset_buffer(self.h, self._stream()))
