#!/usr/bin/env python
"""Times exo_policy_forward on N observations (CUDA events on the launching stream, L2 flushed between launches) and,
beside it, the same networks in plain torch (fp32 matmul and allow_tf32) - the library baseline the fused kernel replaces.
    python tools/policy_probe.py [N] [reps]
"""
import json
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import exo_b200  # noqa: E402

FLOP_PER_OBS = 2 * (80 * 300 * 2 + 300 * 300 * 3 + 600 * 300 + 300 * 7)


def torch_policy(g, dev):
    W = {k: torch.from_numpy(g[k]).to(dev) for k in g.files if "." in k}

    def avg(x):
        return x / x.abs().mean(-1, keepdim=True).clamp(min=1e-8)

    def f(s):
        F = torch.nn.functional
        h = F.elu(F.linear(s, W["zs1.weight"], W["zs1.bias"]))
        h = F.elu(F.linear(h, W["zs2.weight"], W["zs2.bias"]))
        zs = avg(F.linear(h, W["zs3.weight"], W["zs3.bias"]))
        a = avg(F.linear(s, W["l0.weight"], W["l0.bias"]))
        a = torch.relu(F.linear(torch.cat([a, zs], 1), W["l1.weight"], W["l1.bias"]))
        a = torch.relu(F.linear(a, W["l2.weight"], W["l2.bias"]))
        return torch.tanh(F.linear(a, W["l3.weight"], W["l3.bias"])).clamp(-1, 1)
    return f


def timeit(fn, reps, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.fill_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.mean(ts)), float(np.min(ts))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    g = np.load(os.path.join(REPO, "tests", "golden", "ref_policy.npz"))
    dev = torch.device("cuda:0")
    p = exo_b200.DevicePolicy(dev)
    sd = {k: g[k] for k in g.files if "." in k}
    p.set_networks(sd, sd, "checkpoint")
    base = torch.from_numpy(g["obs"]).to(dev)
    obs = base[torch.arange(n, device=dev) % base.shape[0]].contiguous()
    out = torch.empty(n, 7, device=dev)
    flush = torch.empty(512 << 18, device=dev)
    res = {"n": n, "flop_per_obs": FLOP_PER_OBS}
    m, b = timeit(lambda: p.forward(obs, use_checkpoint=True, out=out), reps, flush)
    res["kernel_ms"] = m; res["kernel_ms_min"] = b; res["kernel_tflops"] = n * FLOP_PER_OBS / m / 1e9; res["kernel_obs_per_s"] = n / m * 1e3
    m, b = timeit(lambda: p.forward(obs, use_checkpoint=True, out=out, precision="fp32"), max(3, reps // 3), flush)
    res["fp32_path_ms"] = m; res["fp32_path_tflops"] = n * FLOP_PER_OBS / m / 1e9
    import ctypes as C
    clk = (C.c_int64 * 32)()
    exo_b200._capi.check(p.lib.exo_policy_profile(p._nets["checkpoint"].h, C.c_void_p(obs.data_ptr()), n, C.c_void_p(out.data_ptr()), clk,
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    c = np.array(list(clk), dtype=np.int64)
    t0 = c[16]
    res["timeline_cycles"] = {"mma_phase_start": (c[0:16:2] - t0).tolist(), "mma_phase_issued": (c[1:16:2] - t0).tolist(),
                              "epi_begin": (c[16:32:2] - t0).tolist(), "epi_end": (c[17:32:2] - t0).tolist()}
    tp = torch_policy(g, dev)
    for name, flag in (("torch_fp32", False), ("torch_tf32", True)):
        torch.backends.cuda.matmul.allow_tf32 = flag
        with torch.no_grad():
            m, b = timeit(lambda: tp(obs), reps, flush)
        res[name + "_ms"] = m; res[name + "_tflops"] = n * FLOP_PER_OBS / m / 1e9
    print(json.dumps(res))


if __name__ == "__main__":
    main()
