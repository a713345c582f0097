"""GPU: the general Bullet multibody step in the product (ExoConfig.motor_model = EXO_MOTOR_FULL, csrc/exo_motor.cuh) against
its CPU restatement (oracle/multibody.py, the model behind minibullet's motor_model="full") and against goldens the
reference's own Python env produced on that model (tests/golden/ref_full_motor.npz, oracle/gen_golden.py family F).
Reference call sites: Environment/Exoskeleton_env.py:427-433, Environment/Exoskeleton_sim_pybullet.py:109-117.

Tolerances (stated where used): one multibody step from identical state and targets, fp64: 1e-10 on q, qd and the
prismatic positions (different but equivalent factorisations of M; the sweep count must agree). Closed loop against the
goldens: the goldens integrate the arm ODE with scipy's RK45 (rtol 1e-3) where the kernel uses the exact propagator, and the
solver's early exit is a discontinuity (one sweep more or less moves q by up to the residual, ~1e-5 rad), so observations
are held to 1e-4 (fp64) / 2e-4 (fp32), joint angles to 1e-4 rad."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, TRAIN, TYPES15, case_settings

pytestmark = pytest.mark.gpu

import minibullet
import multibody

# DoF order of the kernel (include/exo_env.h EXO_FIELD_QD): 5 arm joints, humerus dummies act11, act21, act32, act42,
# act52, act62, act72, forearm dummies act12, act22 -> URDF link handles (Exoskeleton_sim_pybullet.py:48-61)
DOF_LINK = [0, 1, 2, 3, 4, 9, 12, 8, 11, 7, 13, 10, 5, 6]
OPTION_SETS = [
    ({}, {}),
    (dict(inertia_from_file=1), dict(inertia_from="file")),
    (dict(row_order=1), dict(row_order="natural")),
    (dict(residual_threshold=0.0, iterations=70), dict(residual_threshold=0.0, iterations=70)),
    (dict(linear_damping=0.0, angular_damping=0.0, kd=0.8), dict(linear_damping=0.0, angular_damping=0.0)),
    (dict(at_limits=True), {}),             # joints at / beyond URDF bounds: joint-limit rows join the solve (0.8 % of bench env-steps)
]


@pytest.mark.parametrize("which", range(len(OPTION_SETS)))
def test_one_multibody_step_matches_the_cpu_model(exo, which):
    kopts, popts = OPTION_SETS[which]
    kopts = dict(kopts)
    at_limits = kopts.pop("at_limits", False)
    geo = dict(np.load(os.path.join(GOLDEN, "ref_data_and_units.npz")))
    n = 10
    mids = [i % 8 for i in range(n)]
    seqs = [TYPES15[(3 * i + 14) % 15] for i in range(n)]
    env = exo.BatchedExoskeletonEnv(mids, seqs, list(range(n)), precision="fp64", motor_model="full", motor_options=kopts, **TRAIN)
    env.reset()
    rs = np.random.RandomState(11 + which)
    # start away from rest so Coriolis / damping / prismatic offsets all matter
    q0 = rs.uniform(-0.4, 0.4, (n, 5)); q0[:, 3] = np.abs(q0[:, 3]) + 0.2
    if at_limits:
        q0[0::3, 3] = -0.05                       # elbow y below its -2 deg lower bound (exo_v3.urdf:77)
        q0[1::3, 0] = 1.45                        # shoulder z above its +80 deg upper bound (:17)
        q0[2::3, 4] = -1.5184364318848            # elbow z exactly on its lower bound: a row with zero penetration
    env.set_field("q", torch.from_numpy(q0).cuda()); env.set_field("q_lag", torch.from_numpy(q0).cuda())
    env.set_field("qd", torch.from_numpy(rs.uniform(-1.0, 1.0, (n, 14)) * np.r_[np.ones(5), 0.05 * np.ones(9)]).cuda())
    env.set_field("q_pris", torch.from_numpy(rs.uniform(-2e-3, 2e-3, (n, 9))).cuda())
    bodies = [minibullet.Body.from_arrays(geo) for _ in range(n)]
    kd = kopts.get("kd", 1.0)
    steppers = [multibody.FullStep(b, 1.0 / 40.0, (0.0, 0.0, -9.81), 0.1, kd, **popts) for b in bodies]
    worst = dict(q=0.0, qd=0.0, qp=0.0)
    sweeps_cpu = 0
    env.stats(reset=True)
    for t in range(12):
        q = env.get_field("q").cpu().numpy(); qd = env.get_field("qd").cpu().numpy(); qp = env.get_field("q_pris").cpu().numpy()
        cnt = env.get_field("counts").cpu().numpy().ravel()
        a = torch.from_numpy(rs.uniform(-1, 1, (n, 7)).astype(np.float32)).cuda()
        _, _, _, info = env.step(a, with_info=True)
        qa = info["ampl_val"].cpu().numpy()
        q1 = env.get_field("q").cpu().numpy(); qd1 = env.get_field("qd").cpu().numpy(); qp1 = env.get_field("q_pris").cpu().numpy()
        for i in range(n):
            m = env.motions[mids[i]][:, cnt[i]]                     # elbow_y, elbow_z, shoulder_x, shoulder_y, shoulder_z (deg)
            tgt = np.array([m[4] + qa[i, 2], m[3] + qa[i, 0], m[2] + qa[i, 1], m[0] + qa[i, 3], m[1]]) * 0.017453292519943295
            b = bodies[i]
            b.q[:] = 0.0; b.qd[:] = 0.0
            b.q[DOF_LINK[:5]] = q[i]; b.q[DOF_LINK[5:]] = qp[i]; b.qd[DOF_LINK] = qd[i]
            b.targets = {j: tgt[j] for j in range(5)}
            before = steppers[i].stats["sweeps"]
            steppers[i].step()
            sweeps_cpu += steppers[i].stats["sweeps"] - before
            worst["q"] = max(worst["q"], np.max(np.abs(b.q[DOF_LINK[:5]] - q1[i])))
            worst["qp"] = max(worst["qp"], np.max(np.abs(b.q[DOF_LINK[5:]] - qp1[i])))
            worst["qd"] = max(worst["qd"], np.max(np.abs(b.qd[DOF_LINK] - qd1[i])))
            assert np.max(np.abs(b.q[[14, 15, 16, 17, 18]])) < 1e-15      # the base dummies the kernel leaves out stay at rest
    st = env.stats()
    limit_rows = sum(sp.stats["limit_rows"] for sp in steppers)
    print("options %r%s: max |dq| %.2e  |dqd| %.2e  |dq_pris| %.2e; sweeps kernel %d cpu %d; limit rows %d" %
          (kopts, " at limits" if at_limits else "", worst["q"], worst["qd"], worst["qp"], int(st[14]), sweeps_cpu, limit_rows))
    assert (limit_rows > 0) == at_limits
    assert worst["q"] <= 1e-10 and worst["qd"] <= 1e-9 and worst["qp"] <= 1e-10
    assert int(st[14]) == sweeps_cpu and int(st[15]) == 12 * n
    env.close()


def _replay(exo, g, cases, precision, settings_case, ode="propagator"):
    mids = [int(g[c + "/meta_motion"]) for c in cases]
    seqs = [g[c + "/meta_seq"] for c in cases]
    seeds = [int(g[c + "/meta_seed"]) for c in cases]
    env = exo.BatchedExoskeletonEnv(mids, seqs, seeds, precision=precision, auto_reset=True, motor_model="full", ode=ode,
                                    **case_settings(g, settings_case))
    obs = env.reset().cpu().numpy()
    nsteps = [len(g[c + "/action"]) for c in cases]
    ep = [0] * len(cases)
    for i, c in enumerate(cases):
        assert np.max(np.abs(obs[i] - g[c + "/ep_obs0"][0])) <= 2e-7
    w = dict(obs=0.0, q=0.0, qd=0.0, qp=0.0, obs0=0.0)
    for t in range(max(nsteps)):
        a = np.stack([g[c + "/action"][min(t, nsteps[i] - 1)] for i, c in enumerate(cases)]).astype(np.float32)
        ob, rw, dn = env.step(torch.from_numpy(a).cuda())
        ob, dn = ob.cpu().numpy(), dn.cpu().numpy()
        q = env.get_field("q").cpu().numpy(); qd = env.get_field("qd").cpu().numpy(); qp = env.get_field("q_pris").cpu().numpy()
        for i, c in enumerate(cases):
            if t >= nsteps[i]:
                continue
            assert bool(dn[i]) == bool(g[c + "/done"][t])
            w["q"] = max(w["q"], float(np.max(np.abs(q[i] - g[c + "/qb"][t]))))
            w["qd"] = max(w["qd"], float(np.max(np.abs(qd[i] - g[c + "/qd19"][t][DOF_LINK]))))
            w["qp"] = max(w["qp"], float(np.max(np.abs(qp[i] - g[c + "/q19"][t][DOF_LINK[5:]]))))
            if dn[i]:
                ep[i] += 1
                w["obs0"] = max(w["obs0"], float(np.max(np.abs(ob[i] - g[c + "/ep_obs0"][ep[i]]))))
            else:
                w["obs"] = max(w["obs"], float(np.max(np.abs(ob[i] - g[c + "/obs"][t]))))
    st = env.stats()
    env.close()
    return w, float(st[14]) / max(float(st[15]), 1.0)


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_closed_loop_against_reference_env_on_the_general_step(exo, precision):
    """The reference's Python env on minibullet(full) vs the CUDA path with EXO_MOTOR_FULL: config 1 with the shipped actor
    over a reset (where the targets jump and the 50-sweep solve is far from converged) and two config-3 episodes."""
    g = np.load(os.path.join(GOLDEN, "ref_full_motor.npz"))
    tol = 1e-4 if precision == "fp64" else 2e-4
    for cases, sc in ((["F0"], "F0"), (["F1", "F2"], "F1")):
        w, sweeps = _replay(exo, g, cases, precision, sc)
        print("%s %s: max |d obs| %.2e (first obs after reset %.2e)  |dq| %.2e  |dqd| %.2e  |dq_pris| %.2e  %.1f sweeps/step"
              % (precision, cases, w["obs"], w["obs0"], w["q"], w["qd"], w["qp"], sweeps))
        assert w["obs"] <= tol and w["obs0"] <= tol and w["q"] <= tol and w["qp"] <= tol, w


def test_closed_loop_general_step_with_the_reference_integrator(exo):
    """EXO_MOTOR_FULL + EXO_ODE_RK45 (fp64): both the integrator and the multibody step are now the ones the golden's
    reference env ran on, so the closed loop must follow it far below the 1e-4 of the test above: the whole state - arm and
    dummy joint positions and velocities of the general step included - to 1e-9, observations to float32 resolution."""
    g = np.load(os.path.join(GOLDEN, "ref_full_motor.npz"))
    for cases, sc in ((["F0"], "F0"), (["F1", "F2"], "F1")):
        w, sweeps = _replay(exo, g, cases, "fp64", sc, ode="rk45")
        print("fp64 rk45 %s: max |d obs| %.2e (first obs after reset %.2e)  |dq| %.2e  |dqd| %.2e  |dq_pris| %.2e  %.1f sweeps/step"
              % (cases, w["obs"], w["obs0"], w["q"], w["qd"], w["qp"], sweeps))
        assert w["obs"] <= 2.4e-7 and w["obs0"] <= 2.4e-7 and w["q"] <= 1e-9 and w["qd"] <= 1e-8 and w["qp"] <= 1e-9, w


def test_general_step_converges_to_the_default_closed_form(exo):
    """With the solver run to convergence the EXO_MOTOR_FULL path and the default EXO_MOTOR_IDEAL path are the same env."""
    n = 16
    mids = [i % 8 for i in range(n)]
    seqs = [TYPES15[i % 15] for i in range(n)]
    kw = dict(precision="fp64", **TRAIN)
    a_env = exo.BatchedExoskeletonEnv(mids, seqs, list(range(n)), **kw)
    b_env = exo.BatchedExoskeletonEnv(mids, seqs, list(range(n)), motor_model="full",
                                      motor_options=dict(iterations=6000, residual_threshold=0.0), **kw)
    assert torch.equal(a_env.reset(), b_env.reset())
    rs = np.random.RandomState(5)
    worst = 0.0
    for t in range(20):
        a = torch.from_numpy(rs.uniform(-1, 1, (n, 7)).astype(np.float32)).cuda()
        oa, ra, _ = a_env.step(a)
        ob, rb, _ = b_env.step(a)
        worst = max(worst, float((oa - ob).abs().max()), float((ra - rb).abs().max()))
    print("converged general step vs closed form, 20 steps: max |d obs|, |d reward| %.2e" % worst)
    assert worst <= 1e-9
    a_env.close(); b_env.close()
