"""CPU: the plain-C oracle (oracle/exo_oracle.c) against the goldens that oracle/gen_golden.py produced by running
the reference's own Python. Everything RNG-derived must be bit-exact; everything else within fp64 round-off."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, case_settings

INFO = ["actuator_torques", "torque_val", "ampl_val", "tremor_torque_val", "tremor_ampl_val"]
REW = ["reward_unwanted", "reward_torque", "reward_axis", "reward_control", "reward_smoothness"]


def test_mt19937_matches_numpy_legacy_stream(oracle_mod):
    for seed in (0, 1, 123, 2 ** 32 - 1):
        np.random.seed(seed)
        assert np.array_equal(np.random.random_sample(2000), oracle_mod.mt_random_samples(seed, 2000))


def test_rk45_matches_scipy_solve_ivp(oracle_mod):
    d = oracle_mod.data()
    for T, q, nfev in zip(d["ode/T"], d["ode/q"], d["ode/nfev"]):
        got, n = oracle_mod.solve_diff_eq_rk45(d["mat/I"], d["mat/D"], d["mat/S"], T)
        assert n == nfev                                  # same accepted/rejected step sequence as scipy
        assert np.max(np.abs(got - q)) <= 1e-13 * max(1.0, np.max(np.abs(q)))


def test_propagator_matches_scipy_expm(oracle_mod):
    import scipy.linalg as sl
    d = oracle_mod.data()
    rs = np.random.RandomState(3)
    for trial in range(6):
        def noisy(M):
            n = rs.uniform(-0.125, 0.125, (7, 7))
            return M + (n + n.T) / 2 * M
        I, D, S = (noisy(d["mat/" + k]) if trial else d["mat/" + k] for k in "IDS")
        Ii = np.linalg.inv(I)
        M = np.zeros((21, 21))
        M[:7, 7:14] = np.eye(7)
        M[7:14, :7], M[7:14, 7:14], M[7:14, 14:] = -Ii @ S, -Ii @ D, Ii
        G = sl.expm(M / 40)[:7, 14:]
        assert np.max(np.abs(oracle_mod.propagator(I, D, S) - G)) <= 1e-11 * np.max(np.abs(G))


def test_exact_solution_vs_rk45_gap_is_the_integrators_tolerance(oracle_mod):
    """Documents why the propagator kernel is compared to the RK45 reference at 1e-3 relative on ampl_val."""
    d = oracle_mod.data()
    G = oracle_mod.propagator(d["mat/I"], d["mat/D"], d["mat/S"])
    worst = max(np.max(np.abs(G @ T - q)) / np.max(np.abs(q)) for T, q in zip(d["ode/T"], d["ode/q"]) if np.any(T))
    assert 1e-7 < worst < 1e-3


def test_forward_kinematics_known_answers(oracle_mod):
    d = oracle_mod.data()
    for th, p in zip(d["fk/theta"], d["fk/p"]):
        assert np.allclose(oracle_mod.forward_kinematics(th), p, atol=1e-15)
    assert np.allclose(oracle_mod.forward_kinematics(np.zeros(7)), [0, 0, -0.8], atol=1e-15)   # SURVEY 8.3 known answer


def _replay(O, g, case, ode_mode=0):
    m = lambda k: g["%s/meta_%s" % (case, k)]   # noqa: E731
    env = O.OracleEnv(int(m("motion")), m("seq"), int(m("seed")), ode_mode=ode_mode, **case_settings(g, case))
    obs, score = env.reset()
    assert score == 2
    ep = 0

    def check_episode(ep, obs):
        E = env.episode()
        for k in ("tt", "I", "D", "S", "shift", "fmax"):
            assert np.array_equal(E[k], g["%s/ep_%s" % (case, k)][ep]), (case, ep, k)     # bit-exact
        assert E["mag"] == float(g[case + "/ep_mag"][ep])
        assert np.array_equal(obs, g[case + "/ep_obs0"][ep])
    check_episode(0, obs)
    A = g[case + "/action"]
    for t in range(len(A)):
        obs, r, dn, trunc, info = env.step(A[t])
        assert trunc is False and dn == bool(g[case + "/done"][t])
        assert np.array_equal(obs, g[case + "/obs"][t]), (case, t)                        # float32-identical
        assert abs(r - g[case + "/reward"][t]) <= 1e-12
        for k in INFO:
            ref = g["%s/%s" % (case, k)][t]
            assert np.max(np.abs(info[k] - ref)) <= 1e-10 * max(1.0, np.max(np.abs(ref))), (case, t, k)
        assert np.max(np.abs(np.array([info[k] for k in REW]) - g[case + "/rew_parts"][t])) <= 1e-12
        assert np.max(np.abs(env.bullet_q() - g[case + "/qb"][t])) <= 1e-13
        if dn:
            obs, _ = env.reset()
            ep += 1
            check_episode(ep, obs)
    return env


def test_cfg1_trained_actor_rollout(oracle_mod):
    g = np.load(os.path.join(GOLDEN, "ref_cfg1_trained_actor.npz"))
    env = _replay(oracle_mod, g, "A0")
    assert env.limit_crossings == 0


@pytest.mark.parametrize("block", range(6))
def test_cfg2_all_types_and_motions(oracle_mod, block):
    g = np.load(os.path.join(GOLDEN, "ref_cfg2_types_x_motions.npz"))
    for case in range(block * 20, block * 20 + 20):
        _replay(oracle_mod, g, "B%d" % case)


@pytest.mark.parametrize("motion", range(8))
def test_cfg3_domain_randomised_episode_and_reset_carryover(oracle_mod, motion):
    g = np.load(os.path.join(GOLDEN, "ref_cfg3_dr_episodes.npz"))
    _replay(oracle_mod, g, "C%d" % motion)


def test_threaded_batch_equals_serial(oracle_mod):
    from conftest import TRAIN, TYPES15
    n = 24
    mids, seqs, seeds = [i % 8 for i in range(n)], [TYPES15[i % 15] for i in range(n)], list(range(n))
    a = oracle_mod.OracleBatch(mids, seqs, seeds, **TRAIN)
    b = oracle_mod.OracleBatch(mids, seqs, seeds, **TRAIN)
    a.reset(); b.reset()
    rs = np.random.RandomState(0)
    for t in range(5):
        act = rs.uniform(-1, 1, (n, 7))
        oa, ra, da = a.step(act, threads=1)
        ob, rb, db = b.step(act, threads=4)
        assert np.array_equal(oa, ob) and np.array_equal(ra, rb) and np.array_equal(da, db)
