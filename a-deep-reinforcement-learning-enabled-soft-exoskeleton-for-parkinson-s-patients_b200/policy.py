"""Host-side mirror of the network part of the reference's Agent.select_action, backed by the fused tcgen05 kernel of
libexo_env.so (include/exo_policy.h, csrc/exo_policy.cu).

  DevicePolicy            holds the (actor, fixed_encoder) and (checkpoint_actor, checkpoint_encoder) pairs the way
                          Agent does (Agent/TD7_multi_agent_Pink_noise.py:151-163); load(filename) reads the files
                          Agent.save writes (:348-366: <filename>_actor, _encoder, _checkpoint_actor, _checkpoint_encoder);
                          forward(obs) takes the batched env's device tensor [N,80] and returns device actions [N,7]
                          with no host copy and no synchronisation; select_action(...) has the reference's signature and
                          numpy-in / numpy-out behaviour (:208-228).
  load_reference_policy   builds one from a directory of the reference's shipped AGENT_NNS checkpoints.

PyTorch supplies device memory and streams (and unpickles the state dicts); every matrix product, normalisation and
activation runs in the CUDA kernel. There is no CPU path.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _capi

STATE_DIM, HIDDEN_DIM, ACTION_DIM = 80, 300, 7
_ENC_KEYS = ("zs1", "zs2", "zs3")
_ACT_KEYS = ("l0", "l1", "l2", "l3")


def _as_f32(x):
    if isinstance(x, torch.Tensor):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(np.asarray(x, dtype=np.float32))


class _Net(object):
    """One (actor, encoder) pair resident on the device: an ExoPolicy handle."""

    def __init__(self, lib, device_index):
        self.lib = lib
        self.h = C.c_void_p()
        _capi.check(lib.exo_policy_create(device_index, C.byref(self.h)))

    def set_weights(self, actor_sd, encoder_sd):
        w = _capi.ExoPolicyWeights()
        keep = []
        for name, sd in [(k, encoder_sd) for k in _ENC_KEYS] + [(k, actor_sd) for k in _ACT_KEYS]:
            for suffix, field in (("weight", "_w"), ("bias", "_b")):
                key = "%s.%s" % (name, suffix)
                if key not in sd:
                    raise KeyError("state dict has no %r (Actor/Encoder of TD7_multi_agent_Pink_noise.py:66-99)" % key)
                arr = _as_f32(sd[key])
                keep.append(arr)
                setattr(w, name + field, arr.ctypes.data_as(C.POINTER(C.c_float)))
        shapes = {"zs1_w": (HIDDEN_DIM, STATE_DIM), "zs2_w": (HIDDEN_DIM, HIDDEN_DIM), "zs3_w": (HIDDEN_DIM, HIDDEN_DIM),
                  "l0_w": (HIDDEN_DIM, STATE_DIM), "l1_w": (HIDDEN_DIM, 2 * HIDDEN_DIM), "l2_w": (HIDDEN_DIM, HIDDEN_DIM),
                  "l3_w": (ACTION_DIM, HIDDEN_DIM)}
        got = dict(zip([n + f for n in _ENC_KEYS + _ACT_KEYS for f in ("_w", "_b")], keep))
        for k, shp in shapes.items():
            if got[k].shape != shp:
                raise _capi.ExoError("%s has shape %s; the kernel is built for %s (state 80 / hidden 300 / action 7)" % (k, got[k].shape, shp))
            if got[k[:-2] + "_b"].shape != (shp[0],):
                raise _capi.ExoError("%s bias has shape %s, expected (%d,)" % (k[:-2], got[k[:-2] + "_b"].shape, shp[0]))
        _capi.check(self.lib.exo_policy_set_weights(self.h, C.byref(w), STATE_DIM, HIDDEN_DIM, ACTION_DIM))

    def close(self):
        if self.h:
            self.lib.exo_policy_destroy(self.h)
            self.h = C.c_void_p()


class DevicePolicy(object):
    """select_action's networks on one B200. `max_action` as Agent.max_action."""

    def __init__(self, device="cuda:0", max_action=1.0, precision="tf32"):
        """precision: "tf32" = the tensor-core kernel (TF32 products, float32 accumulation; torch's allow_tf32 arithmetic),
        "fp32" = float32 products on the FP32 pipe (torch's default matmul; reproduces the reference's actions to
        summation order, ~8x slower). forward(..., precision=...) overrides it per call."""
        if precision not in ("tf32", "fp32"):
            raise ValueError("precision must be 'tf32' or 'fp32'")
        self.precision = precision
        self.lib = _capi.load()
        if not torch.cuda.is_available():
            raise _capi.ExoError("no CUDA device: the policy runs on B200 tensor cores (there is no CPU fallback)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _capi.ExoError("device must be a CUDA device, got %s" % device)
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        self.max_action = float(max_action)
        with torch.cuda.device(idx):
            self._nets = {"current": _Net(self.lib, idx), "checkpoint": _Net(self.lib, idx)}
        self._loaded = set()
        self.noise = None             # [action_dim, ep_length] exploration table (Agent.init_episode_noise, :203-206)

    # ------------------------------------------------------------------ weights
    def set_networks(self, actor_state_dict, encoder_state_dict, which="current"):
        """which: "current" = (actor, fixed_encoder), "checkpoint" = (checkpoint_actor, checkpoint_encoder)."""
        with torch.cuda.device(self.device):
            self._nets[which].set_weights(actor_state_dict, encoder_state_dict)
        self._loaded.add(which)

    def load(self, filename):
        """Agent.load (:349-366) for the four files select_action depends on."""
        def rd(suffix):
            return torch.load(filename + suffix, map_location="cpu", weights_only=False)
        self.set_networks(rd("_actor"), rd("_encoder"), "current")
        self.set_networks(rd("_checkpoint_actor"), rd("_checkpoint_encoder"), "checkpoint")

    # ------------------------------------------------------------------ forward
    def forward(self, obs, use_checkpoint=False, out=None, return_zs=False, precision=None):
        """obs: float32 CUDA tensor [N,80] (e.g. BatchedExoskeletonEnv.step's obs). Returns actions [N,7] on the device,
        clip(actor(obs, zs), -1, 1) * max_action; with return_zs also encoder.zs(obs) [N,300]."""
        which = "checkpoint" if use_checkpoint else "current"
        if which not in self._loaded:
            raise _capi.ExoError("no %s networks loaded" % which)
        if obs.device != self.device or obs.dtype != torch.float32 or obs.dim() != 2 or obs.shape[1] != STATE_DIM:
            raise _capi.ExoError("obs must be a float32 [N,80] tensor on %s" % self.device)
        obs = obs.contiguous()
        n = obs.shape[0]
        if out is None:
            out = torch.empty(n, ACTION_DIM, device=self.device, dtype=torch.float32)
        zs = torch.empty(n, HIDDEN_DIM, device=self.device, dtype=torch.float32) if return_zs else None
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        fn = self.lib.exo_policy_forward if (precision or self.precision) == "tf32" else self.lib.exo_policy_forward_fp32
        _capi.check(fn(self._nets[which].h, C.c_void_p(obs.data_ptr()), n, self.max_action,
                    C.c_void_p(out.data_ptr()), C.c_void_p(zs.data_ptr()) if return_zs else None, stream))
        return (out, zs) if return_zs else out

    __call__ = forward

    def select_action(self, state, timestep=None, first_step=True, use_checkpoint=False, use_exploration=True):
        """Agent.select_action (:208-228): state numpy [B,80] -> numpy [B,7]. With use_exploration the caller provides the
        per-episode pink-noise table in self.noise ([7, ep_length]; generating it is the agent's business, :203-206)."""
        s = torch.from_numpy(np.asarray(state)).float().to(self.device)
        if s.dim() == 1:
            s = s[None]
        prev, self.max_action = self.max_action, 1.0
        try:
            action = self.forward(s, use_checkpoint=use_checkpoint).cpu().numpy()
        finally:
            self.max_action = prev
        if use_exploration:
            if self.noise is None:
                raise _capi.ExoError("use_exploration=True needs the episode noise table in .noise ([7, ep_length])")
            action = action + np.asarray(self.noise)[:, timestep]
        return action.clip(-1, 1) * self.max_action

    def launch_count(self):
        return sum(int(self.lib.exo_policy_launch_count(n.h)) for n in self._nets.values())

    def close(self):
        for n in getattr(self, "_nets", {}).values():
            n.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def load_reference_policy(agent_dir, tremor_type, device="cuda:0", max_action=1.0):
    """The reference ships one agent per tremor type under Simulation/AGENT_NNS/<type>/<type>_* with <type> written as
    "[a,b,c,d]" (Evaluate_control_performance.py:101). tremor_type: that string or four bits."""
    name = tremor_type if isinstance(tremor_type, str) else "[" + ",".join(str(int(b)) for b in tremor_type) + "]"
    p = DevicePolicy(device, max_action)
    p.load(os.path.join(agent_dir, name, name))
    return p
