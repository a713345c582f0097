#!/usr/bin/env python
"""bench.py — batched env-steps/s of the Exoskeleton env hot path on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs-per-gpu E] [--impl reference]

A "step" is one env.step() of every resident environment (one control period dt = 1/40 s each).
Workload (config.workload): BASELINE config 3 per GPU — 65 536 envs/GPU, training domain randomisation
(Exoskeleton_agent_train.py:30-35,43-44), the 8 bundled reference motions in contiguous blocks, the 15 tremor
types round-robin, auto-reset on, U(-1,1)^7 actions generated on the device before timing. Envs shard over
ranks with no data-path collective (weak scaling); NCCL only all-reduces the episode-statistics vector.

Steady state: before anything is timed the population is DE-PHASED (untimed set-up, `dephase`): 347 pre-roll steps in
which env e is reset at pre-roll step e mod 347, so that every later window of steps - however short - contains its share
of episode ends (n / ~280 per step) instead of none for 229 steps and then 8192-env bursts. The line asserts
episodes_finished > 0. The CPU arms de-phase their sample the same way.

Timing: every step sits between its own pair of CUDA events on the launching stream; between steps a 512 MiB
buffer is overwritten to flush the 126 MB L2 (the per-step working set of 65 536 envs would otherwise stay L2
resident). The timed region is bracketed by barrier + synchronize; the reported time is the max over ranks of the
summed per-step device times.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

TYPES15 = [[(m >> 3) & 1, (m >> 2) & 1, (m >> 1) & 1, m & 1, 0, 0, 0] for m in range(1, 16)]
TRAIN = dict(tremor_amplitude_range=[0.1, 1.0], first_harmonics_interval=[4.0, 6.0], second_harmonics_interval=[8.0, 10.0],
             max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.02, dr_actuator_range=0.03,
             matrix_noise_fraction=0.1)
METRIC = "batched env-steps/sec"
UNIT = "env-steps/s"
# Compulsory HBM traffic of one env-step of exo_step_kernel<float,false,4> (DESIGN.md "Algorithmic bytes"):
# reads 520 B + writes 525 B.
ALGO_BYTES_PER_ENV_STEP = 1045
# dram__bytes_read.sum + dram__bytes_write.sum of one exo_step_kernel<float,false,4> launch over 65 536 envs, from the
# ncu --set full capture named in TRAFFIC_SOURCE (refreshed per round from the final code)
# Under the bench's own flush (zeroing 512 MiB leaves L2 full of DIRTY lines) every line the step kernel allocates pushes a
# dirty flush line out to DRAM: r02O, `ncu --cache-control none` with the flush in front of every launch, 12 launches:
# 34.09 MB read (= 520 B/env) + 38.29 MB written = 72.4 MB per launch. With a cold but CLEAN L2 (the flush followed by a read
# pass; also ncu's default cache control, r02e) the same launch reads 34.1 MB and writes 0.1-0.3 MB: its own 34 MB of
# results are still in L2 when it ends. Either way DRAM moves what the algorithm needs (1045 B/env = 68.5 MB) or less.
STEP_KERNEL_DRAM_BYTES_65536 = 34090000 + 38290000
STEP_KERNEL_DRAM_BYTES_65536_CLEAN_L2 = 34120000 + 127000
TRAFFIC_SOURCE = "profiles/r02O_ncu_step_dirty_l2.csv"
DEPHASE_PERIOD = 347          # longest reference motion: every env is reset once during the pre-roll
MEAN_EPISODE_STEPS = float(np.mean([l - 3 for l in (232, 316, 282, 295, 270, 347, 266, 273)]))


def reset_bytes(table_len_mean=MEAN_EPISODE_STEPS + 3, nax=4, real=4):
    """Compulsory HBM bytes of one env reset: MT19937 key read + written back (2 x 2500), the tremor table
    written (L rows x 4 axes), I/D/S (147 fp64), shifts + force maxima (44 words), G (49 words), first observation (320),
    carried state touched (~48 words)."""
    return 2 * 2500 + table_len_mean * nax * real + 147 * 8 + 44 * real + 49 * real + 320 + 48 * real


def settle(env):
    """env.reset() queues a generation pass for every env on the handle's side streams (~2 ms at 65 536 envs): let it
    finish before a timed section, which would otherwise run beside it."""
    import torch
    env.wait_prefetch()
    torch.cuda.synchronize(env.device)


def in_kernel_span_ms(env, actions, flush, reps=6):
    """First warp entry -> last warp exit of exo_step_kernel from the kernel's own timestamps (exo_debug_timeline: per-warp
    globaltimer at entry + clock64 marks), L2 flushed: what the launch costs once the first instruction runs, i.e. the
    CUDA-event figure minus launch latency, completion and event-pair overhead. Informative only; `roofline` uses events."""
    import ctypes as C
    import torch
    nw = (env.n + 31) // 32
    buf = torch.zeros(nw, 8, dtype=torch.int64, device=env.device)
    spans = []
    for rep in range(reps):
        if flush is not None:
            flush.zero_()
        env.lib.exo_debug_timeline(env.h, C.c_void_p(buf.data_ptr()))
        env.step(actions[rep % actions.shape[0]], keep_idle_obs=False)
        torch.cuda.synchronize(env.device)
        env.lib.exo_debug_timeline(env.h, None)
        tl = buf.cpu().numpy()
        t_entry = (tl[:, 0] - tl[:, 0].min()) * 1e-6                      # ms
        life = (tl[:, 7] - tl[:, 2]) / 1.965e6                             # ms at the 1965 MHz the bench runs at
        spans.append(float((t_entry + life).max()))
    return float(np.median(spans[1:]))


def dephase(env, actions, period=DEPHASE_PERIOD):
    """Untimed set-up: spread the envs over their episodes. env e is reset at pre-roll step e mod period; all envs step."""
    import torch
    n = env.n
    phase = torch.arange(n, device=env.device) % period
    for t in range(period):
        env.reset((phase == t).to(torch.uint8))
        env.step(actions[t % actions.shape[0]], keep_idle_obs=False)
    settle(env)


def population(n_per_gpu, rank, world):
    """This rank's shard of the cfg-3/4 population (exo_b200.sharding.benchmark_population)."""
    from exo_b200 import sharding
    return sharding.benchmark_population(n_per_gpu * world, rank, world)


class ClockSampler(object):
    """nvidia-smi clocks/throttle reasons sampled while the timed region runs (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nme, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_baseline(sample_envs, steps, threads=0, seed0=0):
    """The oracle port (oracle/exo_oracle.c, fp64, scipy-RK45-faithful = the reference's arithmetic) on host threads.
    Returns (env_steps_per_s, threads_used, seconds)."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import oracle as O
    motion, seqs, seeds = population(sample_envs, 0, 1)
    batch = O.OracleBatch(motion, seqs, seeds + seed0, ode_mode=0, **TRAIN)
    batch.reset()
    rs = np.random.RandomState(1)
    acts = rs.uniform(-1, 1, (4, sample_envs, 7))
    nthreads = O.lib().orc_max_threads() if threads <= 0 else threads
    batch.dephase(DEPHASE_PERIOD, threads=nthreads)           # same steady state as the GPU arm: episode ends in every window
    batch.step(acts[0], auto_reset=True, threads=nthreads)
    t0 = time.perf_counter()
    for t in range(steps):
        batch.step(acts[t % 4], auto_reset=True, threads=nthreads)
    dt = time.perf_counter() - t0
    return sample_envs * steps / dt, nthreads, dt


def bench_lap(exo_b200, env, actions, dev, flush):
    """Next row N1: the device-resident replay buffer fed by the env's own output tensors.
    Reference hyper-parameters (TD7_multi_agent.py:15-16 Hyperparameters: batch_size 128, buffer_size 2.5e5; 8 rows).
    Reports add throughput (transitions/s and algorithmic GB/s: 2*80+7+3 words written, 2*80+7+2 read per
    transition), and the latency of one sample() + update_priority() over a full buffer."""
    import torch
    n = env.n
    E, MAX, B = 8, 250000, 128
    buf = exo_b200.LAP(80, 7, dev, E, max_size=MAX, batch_size=B)
    obs = env.reset().clone()
    settle(env)
    evs = []
    steps = 40
    for t in range(steps):
        nobs, rew, done = env.step(actions[t % actions.shape[0]])
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        buf.add_batch(obs, actions[t % actions.shape[0]], nobs, rew, done)
        b.record()
        evs.append((a, b))
        obs = nobs
    torch.cuda.synchronize(dev)
    add_ms = float(np.mean([a.elapsed_time(b) for a, b in evs[5:]]))
    size = buf.size
    u = torch.rand(E * B, device=dev)
    pr = torch.rand(E * B, device=dev) + 0.1
    sev = []
    for t in range(25):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        buf.sample(uniforms=u)
        buf.update_priority(pr)
        b.record()
        sev.append((a, b))
    torch.cuda.synchronize(dev)
    samp_ms = float(np.mean([a.elapsed_time(b) for a, b in sev[5:]]))
    bytes_add = n * ((2 * 80 + 7 + 3) + (2 * 80 + 7 + 2)) * 4
    out = {"rows": E, "max_size": MAX, "batch_size": B, "filled_slots_per_row": int(size),
           "add_transitions_per_s": n / (add_ms * 1e-3), "add_ms": add_ms, "add_gbs": bytes_add / (add_ms * 1e-3) / 1e9,
           "sample_plus_update_ms": samp_ms, "l2": "flushed between calls" if flush is not None else "not flushed"}
    buf.close()
    return out


POLICY_FLOP_PER_OBS = 2 * (80 * 300 * 2 + 300 * 300 * 3 + 600 * 300 + 300 * 7)


def synthetic_policy_state_dict(seed=0):
    """Random-init weights of the reference's Actor + Encoder.zs architecture (TD7_multi_agent_Pink_noise.py:59-105:
    zs1 80->300, zs2, zs3 300->300, l0 80->300, l1 600->300, l2 300->300, l3 300->7), torch.nn.Linear's default init."""
    rs = np.random.RandomState(seed)
    sd = {}
    for name, (o, i) in dict(zs1=(300, 80), zs2=(300, 300), zs3=(300, 300), l0=(300, 80), l1=(300, 600), l2=(300, 300), l3=(7, 300)).items():
        b = 1.0 / np.sqrt(i)
        sd[name + ".weight"] = rs.uniform(-b, b, (o, i)).astype(np.float32)
        sd[name + ".bias"] = rs.uniform(-b, b, o).astype(np.float32)
    return sd


def bench_policy(exo_b200, env, dev, flush, peaks, weights=None):
    """Next row N2: select_action's networks (encoder.zs + actor, TD7_multi_agent_Pink_noise.py:66-105)
    as one tcgen05 kernel reading the env's obs tensor. Weights: the shipped [1,1,1,1] checkpoint (tests/golden/ref_policy.npz).
    Reports the forward pass alone (CUDA events, L2 flushed) against the tensor-core roofline - kind::tf32 issues at half
    the bf16 MMA rate, so peak = MEASURED_PEAKS bf16_tflops / 2 - and the closed loop obs -> policy -> env.step."""
    import torch
    n = env.n
    pol = exo_b200.DevicePolicy(dev)
    if weights:                                # a reference checkpoint prefix, e.g. .../AGENT_NNS/[1,1,1,1]/[1,1,1,1]
        pol.load(weights)
    else:
        sd = synthetic_policy_state_dict()
        pol.set_networks(sd, sd, "checkpoint")
    obs = env.reset().clone()
    settle(env)
    act = torch.empty(n, 7, device=dev)
    for _ in range(3):
        pol.forward(obs, use_checkpoint=True, out=act)
    evs = []
    for _ in range(30):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        pol.forward(obs, use_checkpoint=True, out=act)
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize(dev)
    fwd_ms = float(np.mean([a.elapsed_time(b) for a, b in evs[3:]]))
    # the float32 path (FP32 pipe, the reference's own arithmetic), same call
    evs32 = []
    for _ in range(8):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        pol.forward(obs, use_checkpoint=True, out=act, precision="fp32")
        b.record()
        evs32.append((a, b))
    torch.cuda.synchronize(dev)
    fp32_ms = float(np.mean([a.elapsed_time(b) for a, b in evs32[2:]]))
    # closed loop: the policy's actions drive the env (auto-reset on), no host copy anywhere
    steps = 60
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    a.record()
    for _ in range(steps):
        pol.forward(obs, use_checkpoint=True, out=act)
        obs, rew, done = env.step(act)
    b.record()
    torch.cuda.synchronize(dev)
    loop_ms = a.elapsed_time(b) / steps
    # the acting side of the training loop (Exoskeleton_agent_train.py:123-147), all on the device:
    # select_action -> env.step -> replay_buffer.add, for every env, per step
    buf = exo_b200.LAP(80, 7, dev, 8, max_size=250000, batch_size=128)
    prev = obs
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    a.record()
    for _ in range(steps):
        pol.forward(prev, use_checkpoint=True, out=act)
        obs, rew, done = env.step(act)
        buf.add_batch(prev, act, obs, rew, done)
        prev = obs
    b.record()
    torch.cuda.synchronize(dev)
    act_ms = a.elapsed_time(b) / steps
    buf.close()
    tf = n * POLICY_FLOP_PER_OBS / (fwd_ms * 1e-3) / 1e12
    bf16 = float(peaks.get("bf16_tflops", 2250.0))
    out = {"kernel": "exo_policy_kernel", "observations": n, "forward_ms": fwd_ms, "observations_per_s": n / (fwd_ms * 1e-3),
           "flop_per_observation": POLICY_FLOP_PER_OBS, "dtype": "tf32 products, f32 accumulate",
           "roofline": {"bound": "tensor", "achieved": tf, "peak": bf16 / 2, "unit": "TFLOP/s", "frac": tf / (bf16 / 2),
                        "peak_source": ("MEASURED_PEAKS.json bf16_tflops / 2 (kind::tf32 = half the bf16 MMA rate)" if peaks else
                                        "fallback 2250 / 2")},
           "fp32_path": {"kernel": "exo_policy_fp32_kernel", "forward_ms": fp32_ms, "observations_per_s": n / (fp32_ms * 1e-3),
                         "tflops": n * POLICY_FLOP_PER_OBS / (fp32_ms * 1e-3) / 1e12, "dtype": "f32 products and sums (FP32 pipe)"},
           "closed_loop_env_steps_per_s": n / (loop_ms * 1e-3), "closed_loop_ms_per_step": loop_ms,
           "acting_loop_env_steps_per_s": n / (act_ms * 1e-3), "acting_loop_ms_per_step": act_ms,
           "acting_loop": "policy.forward -> env.step -> LAP.add_batch per step, device resident, no host copy",
           "launches": pol.launch_count(),
           "weights": ("checkpoint " + weights) if weights else "synthetic: random-init Actor + Encoder of the reference's architecture"}
    pol.close()
    return out


def bench_td7(exo_b200, dev):
    """Next row N2, second half: one Agent.train() iteration (TD7_multi_agent_Pink_noise.py:230-312) at the reference's
    sizes - 15 tremor types x batch_size 128 = 1920 rows, state 80 / action 7 / zs 300 / hidden 300-320 - with the
    replay buffer on the device: LAP.sample -> update -> update_priority, no host copy, CUDA events."""
    import torch
    envs, bs = 15, 128
    hp = exo_b200.Hyperparameters(batch_size=bs, buffer_size=20000)
    t = exo_b200.DeviceTD7(80, 7, 1, hp=hp, env_num=envs, device=dev, init_seed=1)
    g = torch.Generator(device=dev).manual_seed(3)
    k = 4096
    for _ in range(4):                                      # 16 384 transitions, spread over the 15 rows
        t.replay_buffer.add_batch(torch.randn(k, 80, device=dev, generator=g), torch.rand(k, 7, device=dev, generator=g) * 2 - 1,
                                  torch.randn(k, 80, device=dev, generator=g), torch.randn(k, device=dev, generator=g),
                                  torch.rand(k, device=dev, generator=g) < 0.003)
    for _ in range(10):
        t.train()
    l0 = t.launch_count()
    iters = 100
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    a.record()
    for _ in range(iters):
        t.train()
    b.record()
    torch.cuda.synchronize(dev)
    ms = a.elapsed_time(b) / iters
    batch = t.replay_buffer.sample()
    a.record()
    for _ in range(iters):
        t.train_on_batch(*batch)
    b.record()
    torch.cuda.synchronize(dev)
    ms_upd = a.elapsed_time(b) / iters
    sc = t.scalars()
    gemm = "gemm_kernel<.,.,TC> (mma.sync, csrc/exo_td7.cu)" if os.environ.get("EXO_TD7_GEMM") == "mma" else \
        "tc5::gemm_tc5_kernel (tcgen05.mma kind::tf32 x3, accumulator in tensor memory, csrc/exo_td7_tc5.cuh)"
    out = {"kernel": gemm, "rows": envs * bs, "train_ms": ms, "train_iterations_per_s": 1e3 / ms,
           "update_only_ms": ms_upd, "transitions_per_s": envs * bs / (ms * 1e-3),
           "dtype": "f32 storage and sums, 3xTF32 tensor-core products (fp32-class; EXO_TD7_GEMM=mma = the mma.sync kernel, EXO_TD7_FFMA=1 = FP32-pipe products)",
           "launches_per_iteration": (t.launch_count() - l0) / (2.0 * iters),
           "what": "LAP.sample + encoder/critic/(actor every 2nd) update + Adam + update_priority per iteration, device resident",
           "losses": {k: sc[k] for k in ("encoder_loss", "critic_loss", "actor_loss")}}
    t.close()
    return out


def bench_large_population(exo_b200, dev, flush, peak, precision, n_large=524288, steps=40):
    """The step kernel's roofline fraction away from the single-wave regime: BASELINE config 4 puts 131 072 - 524 288 envs
    on each GPU (1M envs over 8 / 2 GPUs). Same population recipe, same kernel, timed alone with the L2 flushed."""
    import torch
    motion, seqs, seeds = population(n_large, 0, 1)
    env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device=dev, precision=precision, auto_reset=True, **TRAIN)
    gen = torch.Generator(device=dev).manual_seed(99)
    acts = [torch.rand(n_large, 7, device=dev, generator=gen) * 2 - 1 for _ in range(2)]
    env.reset()
    settle(env)
    for t in range(5):
        env.step(acts[t % 2], keep_idle_obs=False)
    ev = []
    for t in range(steps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        env.step(acts[t % 2], keep_idle_obs=False)
        b.record()
        ev.append((a, b))
    torch.cuda.synchronize(dev)
    kms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    env.close()
    achieved = ALGO_BYTES_PER_ENV_STEP * n_large / (kms * 1e-3) / 1e9
    return {"envs": n_large, "kernel_ms": kms, "achieved": achieved, "unit": "GB/s", "frac": achieved / peak,
            "traffic": 491657728 if (n_large == 524288 and precision == "fp32") else None,
            "traffic_source": "ncu --set full dram__bytes_read.sum (272.8 MB = the 520 B/env compulsory reads) + dram__bytes_write.sum "
                              "(218.8 MB of the 525 B/env written; the rest is still in L2 when the launch ends), profiles/r01o_ncu_full_step_524288.txt",
            "note": "same kernel at config 4's per-GPU population (1M envs over 2 GPUs); 65 536 envs are one wave of 14 warps per SM"}


def bench_reset(env, dev, flush, peak):
    """Episode (re)initialisation by itself: exo_begin_kernel (slot flip + first observation, on the caller's stream) and
    exo_gen_kernel (the next episode's parameters: draws, tremor table, I/D/S noise, propagator - on the side stream, one
    episode ahead, OFF the step's critical path; the timed region here waits for it explicitly). (a) one step's steady-state
    share of the population (n / mean episode length envs, random positions, L2 flushed); (b) every env -> throughput and
    the HBM roofline entry."""
    import torch
    n = env.n
    k = max(1, int(round(n / MEAN_EPISODE_STEPS)))
    g = torch.Generator(device="cpu").manual_seed(0)
    was = env.auto_reset
    env.set_auto_reset(False)

    def timed(mask, reps):
        ts = []
        for _ in range(reps):
            if flush is not None:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            env.reset(mask)
            env.wait_prefetch()
            b.record()
            ts.append((a, b))
        torch.cuda.synchronize(dev)
        return float(np.median([a.elapsed_time(b) for a, b in ts[2:]]))
    m = torch.zeros(n, dtype=torch.uint8)
    m[torch.randperm(n, generator=g)[:k]] = 1
    share_ms = timed(m.to(dev), 16)
    empty_ms = timed(torch.zeros(n, dtype=torch.uint8, device=dev), 10)
    all_ms = timed(None, 6)
    env.set_auto_reset(was)
    by = reset_bytes()
    achieved = by * n / (all_ms * 1e-3) / 1e9
    return {"kernel": "exo_gen_kernel (+ exo_begin_kernel)", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": None, "algorithmic_bytes_per_reset": by, "all_envs_ms": all_ms, "ns_per_reset": all_ms * 1e6 / n,
            "resets_per_step_steady_state": k, "steady_state_share_ms": share_ms, "empty_launch_ms": empty_ms,
            "note": "issue/latency-bound (one warp walks an env's draw chain), not bandwidth-bound; the frac is reported against HBM "
                    "because its compulsory work is byte movement. In the timed step this work runs one episode ahead on a side "
                    "stream, overlapped with the step kernels"}


def bench_cfg4(exo_b200, args, rank, world, dev, flush, peak, barrier):
    """BASELINE config 4 in the driver's own run: 1 048 576 envs split over the ranks (N > 1), de-phased, same timing rules;
    the statistics all-reduce (the path's only collective, every 344 steps - it never fires inside a 20-step window) is timed
    explicitly once."""
    import torch
    import torch.distributed as dist
    total = 1 << 20
    n4 = total // world
    motion, seqs, seeds = population(n4, rank, world)
    env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device=dev, precision=args.precision, auto_reset=True, **TRAIN)
    gen = torch.Generator(device=dev).manual_seed(77 + rank)
    acts = torch.rand(2, n4, 7, device=dev, generator=gen) * 2 - 1
    env.reset()
    dephase(env, acts)
    steps = max(5, min(args.steps, 100))
    for t in range(3):
        env.step(acts[t % 2], keep_idle_obs=False)
    env.stats(reset=True)
    barrier()
    ev = []
    for t in range(steps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        env.step(acts[t % 2], keep_idle_obs=False)
        b.record()
        ev.append((a, b))
    barrier()
    tot = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=dev)
    dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    st = env.stats(reset=True).clone()
    exo_b200.sharding.reduce_stats(st)                      # warm the communicator
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    exo_b200.sharding.reduce_stats(st)
    b.record()
    torch.cuda.synchronize(dev)
    ar = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
    dist.all_reduce(ar, op=dist.ReduceOp.MAX)
    env.set_auto_reset(False)
    env.reset()
    settle(env)
    kev = []
    for t in range(20):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        env.step(acts[t % 2], keep_idle_obs=False)
        b.record()
        kev.append((a, b))
    torch.cuda.synchronize(dev)
    kms = float(np.mean([a.elapsed_time(b) for a, b in kev[2:]]))
    env.close()
    ms = float(tot.item()) / steps
    achieved = ALGO_BYTES_PER_ENV_STEP * n4 / (kms * 1e-3) / 1e9
    return {"workload": "BASELINE cfg4: 1 048 576 envs over %d GPUs, de-phased, auto-reset" % world, "total_envs": total,
            "envs_per_gpu": n4, "steps": steps, "value": total * steps / (float(tot.item()) * 1e-3), "unit": UNIT, "ms_per_step": ms,
            "episodes_finished": float(st[0].item()), "step_kernel_ms": kms, "step_kernel_frac_of_hbm": achieved / peak,
            "stats_allreduce_ms": float(ar.item()), "stats_allreduce": "one sum all-reduce of the 16-double statistics vector (NCCL), "
            "issued every 344 steps in production: %.4f of a step when amortised" % (float(ar.item()) / 344.0 / ms)}


def bench_motor_full(exo_b200, args, dev, flush):
    """The same workload with ExoConfig.motor_model = EXO_MOTOR_FULL: the general Bullet multibody step (forward dynamics +
    50-sweep projected Gauss-Seidel, csrc/exo_motor.cuh) instead of its closed-form fixed point. Compute-bound."""
    import torch
    n = args.envs_per_gpu
    motion, seqs, seeds = population(n, 0, 1)
    env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device=dev, precision=args.precision, auto_reset=True,
                                         motor_model="full", **TRAIN)
    gen = torch.Generator(device=dev).manual_seed(5)
    acts = torch.rand(2, n, 7, device=dev, generator=gen) * 2 - 1
    env.reset()
    dephase(env, acts)
    env.stats(reset=True)
    steps = 20
    ev = []
    for t in range(steps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        env.step(acts[t % 2], keep_idle_obs=False)
        b.record()
        ev.append((a, b))
    torch.cuda.synchronize(dev)
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    st = env.stats().cpu().numpy()
    env.close()
    return {"motor_model": "EXO_MOTOR_FULL (pybullet default solver settings: 50 iterations, residual threshold 1e-7, inertia from "
                           "collision boxes)", "envs": n, "ms_per_step": ms, "value": n / (ms * 1e-3), "unit": UNIT,
            "solver_sweeps_per_env_step": float(st[14] / max(st[15], 1.0)), "kernel": "exo_motor_kernel + exo_step_kernel<FULL>"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU algorithm (oracle port; the Python reference cannot travel to the GPU box
    and needs pybullet) on all host threads, rank 0 only."""
    if rank != 0:
        return
    sample = args.ref_sample
    # warm-up
    cpu_baseline(sample, max(1, min(args.warmup, 20)))
    t0 = time.perf_counter()
    v, cores, secs = cpu_baseline(sample, args.steps)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "cfg3 population, bounded sample of %d envs per step on host threads" % sample,
                       "envs_per_step": sample},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "%d envs x %d steps, oracle/exo_oracle.c (fp64, RK45 as scipy), %d pthreads" % (sample, args.steps, cores)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--envs-per-gpu", type=int, default=65536)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--e2e-steps", type=int, default=30)
    ap.add_argument("--kernel-steps", type=int, default=100)
    ap.add_argument("--cpu-sample", type=int, default=4096, help="envs in the bounded CPU-baseline sample")
    ap.add_argument("--cpu-steps", type=int, default=2000)
    ap.add_argument("--ref-sample", type=int, default=4096)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--no-lap", action="store_true", help="skip the replay-buffer (next row N1) measurement")
    ap.add_argument("--no-large", action="store_true", help="skip the 524 288-env roofline measurement of the step kernel")
    ap.add_argument("--no-policy", action="store_true", help="skip the policy forward pass (next row N2) measurement")
    ap.add_argument("--policy-weights", default=None, help="reference checkpoint prefix for the policy leg (default: synthetic weights)")
    ap.add_argument("--no-dephase", action="store_true", help="skip the pre-roll (lock-step population: no episode ends for 229 steps)")
    ap.add_argument("--no-cfg4", action="store_true", help="skip the BASELINE config 4 sub-record (1M envs over the ranks, N > 1 only)")
    ap.add_argument("--no-motor-full", action="store_true", help="skip the EXO_MOTOR_FULL (general multibody step) sub-record")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    import exo_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    host_pin = exo_b200.hostpin.pin_to_gpu(local_rank)     # before any pinned allocation: the host buffers land next to the GPU
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n = args.envs_per_gpu
    motion, seqs, seeds = population(n, rank, world)
    env = exo_b200.BatchedExoskeletonEnv(motion, seqs, seeds, device=dev, precision=args.precision, auto_reset=True, **TRAIN)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    n_act = 8
    actions = torch.rand(n_act, n, 7, device=dev, generator=gen) * 2 - 1
    flush = None if args.no_flush else torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    stats_sum = torch.zeros(16, dtype=torch.float64, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # nvidia-smi is started ahead of the untimed set-up: its start-up (a process spawn + NVML initialisation per rank) would
    # otherwise compete with the host thread that launches the first timed steps; it keeps sampling through the timed region
    sampler = ClockSampler(local_rank)
    sampler.start()
    env.reset()
    if not args.no_dephase:
        dephase(env, actions)
    for t in range(args.warmup):
        env.step(actions[t % n_act])
    env.stats(reset=True)
    barrier()
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    launches0 = env.launch_count
    wall0 = time.perf_counter()
    for t in range(args.steps):
        if flush is not None:
            flush.zero_()
        ev0[t].record()
        env.step(actions[t % n_act])
        if world > 1 and (t + 1) % 344 == 0:          # episode statistics: the only collective (cfg 4)
            stats_sum += exo_b200.sharding.reduce_stats(env.stats(reset=True).clone())
        ev1[t].record()
    barrier()
    wall = time.perf_counter() - wall0
    launches = env.launch_count - launches0
    per_step = np.array([a.elapsed_time(b) for a, b in zip(ev0, ev1)])
    total_ms = torch.tensor([per_step.sum()], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    stats_sum += exo_b200.sharding.reduce_stats(env.stats(reset=True).clone())
    value = world * n * args.steps / (total_ms * 1e-3)

    # ---- dominant kernel alone (exo_step_kernel): auto-reset off, all envs live, flush between launches
    env.set_auto_reset(False)
    env.reset()
    settle(env)
    ksteps = min(args.kernel_steps, 200)
    for t in range(3):
        env.step(actions[t % n_act], keep_idle_obs=False)
    k0 = [torch.cuda.Event(enable_timing=True) for _ in range(ksteps)]
    k1 = [torch.cuda.Event(enable_timing=True) for _ in range(ksteps)]
    torch.cuda.synchronize(dev)
    for t in range(ksteps):
        if flush is not None:
            flush.zero_()
        k0[t].record()
        env.step(actions[t % n_act], keep_idle_obs=False)
        k1[t].record()
    torch.cuda.synchronize(dev)
    kms = float(np.mean([a.elapsed_time(b) for a, b in zip(k0, k1)]))
    # the same with a cold but clean L2: the flush's dirty lines are drained by a read pass before the launch
    kms_clean = kms
    if flush is not None:
        for t in range(ksteps):
            flush.zero_()
            flush.view(torch.int32).sum()
            k0[t].record()
            env.step(actions[t % n_act], keep_idle_obs=False)
            k1[t].record()
        torch.cuda.synchronize(dev)
        kms_clean = float(np.mean([a.elapsed_time(b) for a, b in zip(k0, k1)]))
    span_ms = in_kernel_span_ms(env, actions, flush) if rank == 0 else None
    clocks = sampler.stop()

    # ---- end to end through the host-buffer C-ABI call (exo_step_host): pinned H2D + kernels + D2H + sync per step.
    #      First the ceiling: the bare pinned device->host copy of one step's observations from every rank at once.
    env.set_auto_reset(True)
    env.reset()
    settle(env)
    obs_h = torch.zeros(n, 80).pin_memory()
    obs_d = torch.zeros(n, 80, device=dev)
    for _ in range(3):
        obs_h.copy_(obs_d, non_blocking=True)
    barrier()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    for _ in range(20):
        obs_h.copy_(obs_d, non_blocking=True)
    c1.record()
    barrier()
    d2h_ms = torch.tensor([c0.elapsed_time(c1) / 20.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(d2h_ms, op=dist.ReduceOp.MAX)
    d2h_ms = float(d2h_ms.item())
    del obs_h, obs_d
    host_actions = [(torch.rand(n, 7) * 2 - 1).pin_memory() for _ in range(4)]
    for t in range(3):
        env.step_host(host_actions[t % 4])
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.e2e_steps)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.e2e_steps)]
    barrier()
    for t in range(args.e2e_steps):
        if flush is not None:
            flush.zero_()
        e0[t].record()
        env.step_host(host_actions[t % 4])
        e1[t].record()
    barrier()
    e2e_ms = torch.tensor([sum(a.elapsed_time(b) for a, b in zip(e0, e1))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = world * n * args.e2e_steps / (float(e2e_ms.item()) * 1e-3)
    rbytes = 4 if args.precision == "fp32" else 8

    lap_line = None
    if rank == 0 and not args.no_lap:
        lap_line = bench_lap(exo_b200, env, actions, dev, flush)
    peaks = {}
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    policy_line = None
    if rank == 0 and not args.no_policy:
        policy_line = bench_policy(exo_b200, env, dev, flush, peaks, args.policy_weights)
        policy_line["td7_update"] = bench_td7(exo_b200, dev)
    peak = float(peaks.get("hbm_gbs", 6650.0))
    cfg4 = None
    if world > 1 and not args.no_cfg4:
        cfg4 = bench_cfg4(exo_b200, args, rank, world, dev, flush, peak, barrier)
    if rank == 0:
        reset_line = bench_reset(env, dev, flush, peak)
        motor_line = None
        if world == 1 and not args.no_motor_full:
            motor_line = bench_motor_full(exo_b200, args, dev, flush)
        achieved = ALGO_BYTES_PER_ENV_STEP * n / (kms * 1e-3) / 1e9
        cpu_v, cpu_cores, cpu_secs = cpu_baseline(args.cpu_sample, args.cpu_steps)
        large = None
        if world == 1 and not args.no_large and args.precision == "fp32":
            large = bench_large_population(exo_b200, dev, flush, peak, args.precision)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if args.precision == "fp32" else "f64", "data": "synthetic",
            "config": {"workload": "BASELINE cfg3 per GPU: %d envs/GPU, training DR (0.02/0.03/0.1), 8 reference motions x 15 tremor "
                                   "types, auto-reset, U(-1,1)^7 device actions, population de-phased by a %d-step untimed pre-roll"
                                   % (n, 0 if args.no_dephase else DEPHASE_PERIOD),
                       "expected_episode_ends_per_step": n / MEAN_EPISODE_STEPS,
                       "envs_per_gpu": n, "total_envs": world * n, "l2": "flushed between steps (512 MiB overwrite, untimed)"
                       if flush is not None else "not flushed", "episodes_finished": stats_sum[0].item(),
                       "mean_return": (stats_sum[1] / stats_sum[0]).item() if stats_sum[0].item() > 0 else None,
                       "wall_s_timed_region": wall},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": STEP_KERNEL_DRAM_BYTES_65536 if (n == 65536 and args.precision == "fp32") else None,
                         "traffic_source": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum of one launch, "
                                           + TRAFFIC_SOURCE + " (--cache-control none, the bench's flush before every launch: 34.1 MB "
                                           "read + 38.3 MB of the flush's dirty lines written back; with a clean cold L2 the launch "
                                           "reads 34.1 MB and writes 0.1 MB, profiles/r02O_ncu_step_clean_l2.csv)",
                         "traffic_clean_l2": STEP_KERNEL_DRAM_BYTES_65536_CLEAN_L2 if (n == 65536 and args.precision == "fp32") else None,
                         "kernel_ms_clean_l2": kms_clean,
                         "clean_l2_note": "same launches with the flush followed by a read pass over the flush buffer (L2 cold, no dirty "
                                          "lines to write back): 1045 B x envs / kernel_ms_clean_l2 = %.0f GB/s = %.2f of peak. Informative; "
                                          "`frac` uses the write-flushed launches." % (
                                              ALGO_BYTES_PER_ENV_STEP * n / (kms_clean * 1e-3) / 1e9, ALGO_BYTES_PER_ENV_STEP * n / (kms_clean * 1e-3) / 1e9 / peak),
                         "kernel": "exo_step_kernel", "kernel_ms": kms,
                         "in_kernel_span_ms": span_ms,
                         "in_kernel_span_note": "first warp entry -> last warp exit from the kernel's own clocks (exo_debug_timeline), "
                                                "instrumented launches: the event figure above minus launch / completion / event-pair "
                                                "overhead; 1045 B x envs / span = %.0f GB/s = %.2f of peak. Not used for `frac`."
                                                % (ALGO_BYTES_PER_ENV_STEP * n / (span_ms * 1e-3) / 1e9, ALGO_BYTES_PER_ENV_STEP * n / (span_ms * 1e-3) / 1e9 / peak),
                         "algorithmic_bytes_per_env_step": ALGO_BYTES_PER_ENV_STEP,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
            "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": cpu_cores, "kind": "port",
                             "sample": "%d envs x %d steps (%.1f s) of the same population, oracle/exo_oracle.c fp64 with scipy-RK45 "
                                       "restated, %d pthreads" % (args.cpu_sample, args.cpu_steps, cpu_secs, cpu_cores)},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * 7 * 4, "d2h_bytes_per_step": n * (80 * 4 + rbytes + 1),
                    "steps": args.e2e_steps, "call": "exo_step_host (pinned host buffers)",
                    "d2h_ceiling": {"what": "bare concurrent pinned D2H of one step's observations (%.1f MB) from all %d ranks" % (n * 320 / 1e6, world),
                                    "gbs_per_rank": n * 320 / (d2h_ms * 1e-3) / 1e9, "env_steps_per_s": world * n / (d2h_ms * 1e-3),
                                    "e2e_fraction_of_ceiling": e2e_value / (world * n / (d2h_ms * 1e-3))},
                    "host_placement": host_pin},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        line["roofline_reset"] = reset_line
        if motor_line is not None:
            line["motor_full"] = motor_line
        if cfg4 is not None:
            line["cfg4"] = cfg4
        if not args.no_dephase and not (stats_sum[0].item() > 0):
            raise SystemExit("bench.py: no episode ended inside the timed region - the steady-state assumption is broken")
        if large is not None:
            line["roofline"]["large_population"] = large
        if lap_line is not None:
            line["lap"] = lap_line
        if policy_line is not None:
            line["policy"] = policy_line
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
