"""CPU: the general Bullet multibody step (oracle/multibody.py, minibullet motor_model="full") - its own consistency,
and how far the closed-form position tracker the product uses by default (q+ = q + kp (target - q)) is from
it on the target sequences of the golden families. Reference call sites: Environment/Exoskeleton_env.py:427-433,
Environment/Exoskeleton_sim_pybullet.py:109-117; tree and inertial data: Simulation/exo_v3.urdf (via
tests/golden/ref_data_and_units.npz geo/*).

The numbers printed here (pytest -s) are the ones DESIGN.md section 4 tabulates; the closed loop version of the same
comparison (reference env on both motor models) is oracle/bullet_model_report.py -> profiles/r02b_bullet_model_report.txt."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

import minibullet
import multibody

DT = 1.0 / 40.0
G = (0.0, 0.0, -9.81)


@pytest.fixture(scope="module")
def geo():
    return dict(np.load(os.path.join(GOLDEN, "ref_data_and_units.npz")))


def test_tree_from_golden_arrays_reproduces_zero_pose(geo):
    b = minibullet.Body.from_arrays(geo)
    assert b.n == 19
    assert np.max(np.abs(np.array(b.cache) - geo["geo/zero_pose_com"])) == 0.0
    assert abs(sum(l.mass for l in b.links) - (0.2 * 2 + 2.0 + 0.1122 + 1.122 + 14.0)) < 1e-6      # exo_v3.urdf masses


@pytest.mark.parametrize("inertia_from", ["collision", "file"])
def test_mass_matrix_and_bias_are_consistent(geo, inertia_from):
    """M symmetric positive definite; with damping off the power balance d/dt (T + V) = 0 holds for qdd = -M^-1 bias
    (checks the Coriolis / centrifugal / gyroscopic terms of the recursive Newton-Euler pass against the composite M)."""
    b = minibullet.Body.from_arrays(geo)
    dyn = multibody.TreeDynamics(b, inertia_from, 0.0, 0.0)
    rs = np.random.RandomState(1)
    for trial in range(4):
        q = rs.uniform(-0.6, 0.6, 19)
        q[5:] *= 0.05
        qd = rs.uniform(-1.5, 1.5, 19)
        kin = dyn.kinematics(q)
        M = dyn.mass_matrix(kin)
        assert np.max(np.abs(M - M.T)) < 1e-14 and np.linalg.eigvalsh(M)[0] > 0
        qdd = -np.linalg.solve(M, dyn.bias(kin, qd, G))

        def energy(q_, qd_):
            k = dyn.kinematics(q_)
            return 0.5 * qd_ @ dyn.mass_matrix(k) @ qd_ + sum(dyn.mass[i] * 9.81 * k["com"][i][2] for i in range(19))
        h = 1e-6
        dE = (energy(q + h * qd, qd + h * qdd) - energy(q - h * qd, qd - h * qdd)) / (2 * h)
        assert abs(dE) < 1e-6 * max(1.0, abs(qd @ dyn.bias(kin, qd, G)))
    # the 5 dummies on the fixed base do not couple with the arm: their rows of M are diagonal (1 kg each)
    base_pris = [i for i, l in enumerate(b.links) if l.parent < 0 and l.joint_type == "prismatic"]
    assert len(base_pris) == 5
    for i in base_pris:
        off = np.delete(M[i], i)
        assert np.max(np.abs(off)) == 0.0 and abs(M[i, i] - 1.0) < 1e-12


def test_bullet_quicksort_row_order_is_a_permutation_with_known_head():
    p = multibody.bullet_quicksort_equal_keys(38)
    assert sorted(p) == list(range(38))
    # first partition pass reverses the array, the recursion then permutes each half again (restated algorithm)
    assert p[:5] == [25, 24, 27, 26, 23] and p[-5:] == [14, 11, 10, 13, 12]
    assert multibody.bullet_quicksort_equal_keys(1) == [0] and multibody.bullet_quicksort_equal_keys(2) == [1, 0]


def _targets_from_golden(qb):
    """The reference's position targets, recovered from the Bullet joint angles the golden recorded after each step
    (the goldens ran on the ideal tracker: q+ = q + 0.1 (t - q))."""
    q_prev = np.vstack([np.zeros((1, 5)), qb[:-1]])
    return q_prev + (qb - q_prev) / 0.1


def _replay(geo, targets, **knobs):
    w = minibullet.World(motor_model="full", **knobs)
    w.bodies.append(minibullet.Body.from_arrays(geo))
    w.full.append(None)
    w.dt, w.gravity = DT, G
    out = []
    for t in targets:
        w.set_position_targets(0, range(5), t)
        w.step_simulation()
        out.append(w.bodies[0].q.copy())
    return np.array(out), w.full[0].stats


def test_closed_form_is_the_converged_solution_of_the_general_step(geo):
    """With the constraint solver run to convergence the general step (gravity, damping, Coriolis, 19 DoF, motor rows
    with Bullet's caps) lands exactly on the closed form: the tracker is not an approximation of the dynamics, only
    of Bullet's 50-sweep truncation of the solve."""
    g = np.load(os.path.join(GOLDEN, "ref_cfg2_types_x_motions.npz"))
    worst = 0.0
    for case in (0, 37, 119):
        qb = g["B%d/qb" % case]
        got, st = _replay(geo, _targets_from_golden(qb), iterations=4000, residual_threshold=0.0)
        worst = max(worst, np.max(np.abs(got[:, :5] - qb)), np.max(np.abs(got[:, 5:])))
        assert st["cap_hits"] == 0 and st["limit_rows"] == 0
    print("converged general step vs closed form: max |dq| %.2e rad (incl. prismatic drift)" % worst)
    assert worst < 1e-10


def test_pybullet_default_solver_settings_vs_closed_form(geo):
    """pybullet's defaults (50 sweeps, early exit at squared residual 1e-7) truncate the solve. Away from episode
    boundaries (family B: 30 steps from the zero pose, targets move <= 2.8 deg/step + tremor) the truncation leaves
    <= 2e-4 rad; the bound below is the measured value plus margin. The large steps are the first of an episode
    (target jumps back to the start of the motion): see the closed-loop report for families A and C."""
    g = np.load(os.path.join(GOLDEN, "ref_cfg2_types_x_motions.npz"))
    rows = []
    for name, knobs in (("collision-box inertias (pybullet default)", {}), ("URDF inertias", dict(inertia_from="file")),
                        ("no early exit", dict(residual_threshold=0.0))):
        dq = pris = 0.0
        sweeps = steps = 0
        for case in range(0, 120, 9):
            qb = g["B%d/qb" % case]
            got, st = _replay(geo, _targets_from_golden(qb), **knobs)
            dq = max(dq, np.max(np.abs(got[:, :5] - qb)))
            pris = max(pris, st["max_prismatic"])
            sweeps += st["sweeps"]; steps += st["steps"]
        rows.append((name, dq, pris, sweeps / steps))
        print("%-45s max |dq| %.2e rad  max prismatic drift %.2e m  %.1f sweeps/step" % rows[-1])
    assert rows[0][1] < 5e-4 and rows[1][1] < 1e-4 and rows[2][1] < 5e-4
    assert all(r[2] < 1e-4 for r in rows)
