"""GPU: exo_policy_forward_fp32_noalloc (include/exo_policy.h) - the allocation-free variant of exo_policy_forward_fp32
whose per-CTA scratch is a static __device__ array in csrc/exo_policy_fp32.cu. No counterpart in the reference,
so the checker is the existing entry point: same kernel, same arithmetic, hence the
tolerance is bit-exact (tests/test_policy.py holds exo_policy_forward_fp32 itself to 1e-5 / 1e-4 against the reference's
torch modules; that bound is repeated here against the golden actions).

Collected last on purpose. Calls go through the C-ABI (ctypes), two calls in a row with
different inputs, so that the second call sees the first call's contents in the shared scratch."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

MAX_OBS = 4096          # EXO_POLICY_NOALLOC_MAX_OBS (include/exo_policy.h)


def test_header_states_the_limit():
    text = open(os.path.join(os.path.dirname(GOLDEN), "..", "include", "exo_policy.h")).read()
    assert "#define EXO_POLICY_NOALLOC_MAX_OBS %d" % MAX_OBS in text


def test_noalloc_forward_equals_the_existing_entry_point_twice_in_a_row(exo):
    import torch
    gold = np.load(os.path.join(GOLDEN, "ref_policy.npz"))
    p = exo.DevicePolicy("cuda:0", max_action=1.0)
    sd = {k: gold[k] for k in gold.files if "." in k}
    p.set_networks(sd, sd, "checkpoint")
    lib, h = p.lib, p._nets["checkpoint"].h
    dev = torch.device("cuda:0")
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    base = torch.from_numpy(gold["obs"]).to(dev)

    def call(fn, obs, with_zs=True):
        n = obs.shape[0]
        act = torch.full((n, 7), float("nan"), device=dev)
        zs = torch.full((n, 300), float("nan"), device=dev) if with_zs else None
        rc = fn(h, C.c_void_p(obs.data_ptr()), n, 1.0, C.c_void_p(act.data_ptr()), C.c_void_p(zs.data_ptr()) if with_zs else None, stream)
        return rc, act, zs

    g = torch.Generator(device="cuda").manual_seed(5)
    # call 1: the golden observations (ragged: not a multiple of the 64-row tile unless the fixture happens to be);
    # call 2, straight after it on the same stream: different rows, the maximum size, scratch still holding call 1's values
    obs1 = base[: min(base.shape[0], MAX_OBS)].contiguous()
    idx = torch.randint(0, base.shape[0], (MAX_OBS,), device=dev, generator=g)
    obs2 = (base[idx] * (1.0 + 0.05 * torch.randn(MAX_OBS, 80, device=dev, generator=g))).contiguous()
    outs = []
    for obs in (obs1, obs2):
        rc, act, zs = call(lib.exo_policy_forward_fp32_noalloc, obs)
        assert rc == 0, lib.exo_last_error().decode()
        outs.append((act, zs))
    torch.cuda.synchronize()
    for obs, (act, zs) in zip((obs1, obs2), outs):
        rc, act_ref, zs_ref = call(lib.exo_policy_forward_fp32, obs)
        assert rc == 0
        torch.cuda.synchronize()
        assert not torch.isnan(act).any() and not torch.isnan(zs).any()
        assert torch.equal(act, act_ref) and torch.equal(zs, zs_ref)            # same kernel, same arithmetic: bit-exact
    # and against the reference's own modules, at the tolerance test_policy.py uses for the fp32 path
    n1 = obs1.shape[0]
    assert np.abs(outs[0][0].cpu().numpy() - gold["action"][:n1]).max() <= 1e-4
    assert np.abs(outs[0][1].cpu().numpy() - gold["zs"][:n1]).max() <= 1e-5
    # ragged sizes, zs omitted
    for n in (1, 63, 65, MAX_OBS - 1):
        rc, act, _ = call(lib.exo_policy_forward_fp32_noalloc, obs2[:n].contiguous(), with_zs=False)
        assert rc == 0
        torch.cuda.synchronize()
        assert torch.equal(act, outs[1][0][:n])
    # beyond the maximum: refused with a message that names the limit, nothing written
    big = torch.zeros(MAX_OBS + 1, 80, device=dev)
    rc, act, _ = call(lib.exo_policy_forward_fp32_noalloc, big, with_zs=False)
    torch.cuda.synchronize()
    assert rc != 0 and torch.isnan(act).all()
    msg = lib.exo_last_error().decode()
    assert str(MAX_OBS) in msg and "EXO_POLICY_NOALLOC_MAX_OBS" in msg
    p.close()
