"""The per-env part of Simulation/Evaluate_control_performance.py:128-330 on the device, and its console layout - the text
the reference keeps under Evaluation_logs/<tremor type> (row N4 of SURVEY.md §8.6: a drop-in for tooling that reads them).

One env per bundled reference motion with the given tremor sequence, the checkpoint actor / encoder pair without
exploration (:157), info per step (env.step(with_info=True)); the script's per-step arithmetic (:187-252: torque and
amplitude suppression percentages with eps = 1e-10 and nan_to_num, the D-H axis swap, forward_kinematics distances,
"all axes" / "any axis" suppression counters, the `disregard` masking) is applied to the recorded [T, n, 7] tensors in one
vectorised pass per episode instead of inside the step loop. Plots, violin / mesh exports and the cross-episode "GLOBAL"
summary of the script are not reproduced.
"""
import io

import numpy as np
import torch

from .env import BatchedExoskeletonEnv
# the evaluation script's own argument defaults (Evaluate_control_performance.py:30-46)
EVAL_DEFAULTS = dict(tremor_amplitude_range=[0.95, 1.05], first_harmonics_interval=[3.75, 6.25], second_harmonics_interval=[7.5, 12.5],
                     max_force_shoulder=40.0, max_force_elbow=20.0, dr_actuator_end_pos_shift=0.025, dr_actuator_range=0.04,
                     matrix_noise_fraction=0.125)


def _fmt(a):
    # the script runs under np.set_printoptions(precision=3, floatmode="fixed"): equal-width columns
    return np.array2string(np.asarray(a, dtype=np.float64), precision=3, floatmode="fixed", suppress_small=True, max_line_width=200)


def evaluate_device(policy, tremor_sequence, num_eval_episodes=1, num_reference_motions=8, device="cuda:0", seed=0, precision="fp64",
                    env_kwargs=None, out=None, humerus_length=0.4, forearm_length=0.4, hand_length=0.05):
    """Runs num_eval_episodes evaluation episodes. Returns (metrics, text): metrics[episode][env] is a dict of the printed
    quantities, text the console output in the script's layout (also written to `out` if given: a path or a file object)."""
    dev = torch.device(device)
    seq = np.asarray(tremor_sequence, np.uint8)
    n = num_reference_motions
    kw = dict(EVAL_DEFAULTS)
    kw.update(env_kwargs or {})
    env = BatchedExoskeletonEnv(np.arange(n, dtype=np.int32), seq, np.arange(n, dtype=np.uint32) + 7919 * int(seed), device=dev,
                                precision=precision, auto_reset=False, **kw)
    max_len = np.asarray(env.max_count, dtype=np.int64)                    # env.return_max_length()
    sel = seq[:7] == 1
    eps = 1e-10
    buf = io.StringIO()
    say = lambda *a: print(*a, file=buf)                                   # noqa: E731
    say("Seed successfully set")
    scores, agent_rew, metrics = [], [], []
    for _ in range(num_eval_episodes):
        obs = env.reset()
        tt = env.get_field("tremor_table")                                 # [n, L, axes]: return_generated_tremor_data (:131)
        torque_max = np.zeros((n, 7))
        torque_max[:, :tt.shape[2]] = np.stack([tt[i, :max_len[i]].max(0).values.cpu().numpy() for i in range(n)])
        rec = {k: [] for k in ("torque_val", "tremor_torque_val", "ampl_val", "tremor_ampl_val", "reward", "live")}
        done_prev = torch.zeros(n, dtype=torch.bool, device=dev)
        for _t in range(int(max_len.max()) + 2):
            act = policy.forward(obs, use_checkpoint=True)
            obs, rew, done, info = env.step(act, with_info=True)
            for k in ("torque_val", "tremor_torque_val", "ampl_val", "tremor_ampl_val"):
                rec[k].append(info[k].double().clone())
            rec["reward"].append(rew.double().clone())
            rec["live"].append(~done_prev)
            done_prev = done_prev | done
            if bool(done_prev.all()):
                break
        T = len(rec["live"])
        live = torch.stack(rec["live"]).cpu().numpy()                      # [T, n]: env i took step t
        R = torch.stack(rec["reward"]).cpu().numpy() * live
        tq, ttq, am, tam = (torch.stack(rec[k]).cpu().numpy() for k in ("torque_val", "tremor_torque_val", "ampl_val", "tremor_ampl_val"))
        with np.errstate(divide="ignore", invalid="ignore"):
            red = np.nan_to_num((np.abs(tq) - np.abs(ttq)) / np.abs(ttq + eps) * 100, nan=0, posinf=0, neginf=0)            # :187-188
            red_ampl = np.nan_to_num((np.abs(am) - np.abs(tam)) / np.abs(tam + eps) * 100, nan=0, posinf=0, neginf=0)       # :191-192
        # total amplitude in the end-effector position (:196-224); counts = 3 + t after step t
        orig = np.zeros((T, n, 7))
        for i in range(n):
            m = env.motions[i]
            c = np.minimum(3 + np.arange(T), m.shape[1] - 1)
            orig[:, i, :5] = np.stack([m[2, c], m[3, c], m[4, c], m[0, c], m[1, c]], 1)
        base = np.radians(orig)
        am_s, tam_s = am.copy(), tam.copy()
        am_s[..., [0, 1]] = am_s[..., [1, 0]]                              # the D-H swap of SAA / SFE (:199-200)
        tam_s[..., [0, 1]] = tam_s[..., [1, 0]]
        fk = lambda th: env.forward_kinematics(torch.from_numpy(th.reshape(-1, 7)), humerus_length, forearm_length,  # noqa: E731
                                               hand_length).double().cpu().numpy().reshape(T, n, 3)
        p0, p1, p2 = fk(base), fk(np.radians(am_s) + base), fk(np.radians(tam_s) + base)
        d_sup, d_unsup = np.linalg.norm(p0 - p1, axis=-1), np.linalg.norm(p0 - p2, axis=-1)
        with np.errstate(divide="ignore", invalid="ignore"):
            red_total = (d_sup - d_unsup) / d_unsup * 100                  # :224
        all_axes = np.all(red[..., sel] <= 0, axis=-1) & live              # :231-236
        any_axes = np.any(red[..., sel] <= 0, axis=-1) & live
        ampl_neg = (red_total < 0) & live                                  # :241-245
        red_total_ep = np.where(ampl_neg, red_total, 0.0)
        red_d, red_ampl_d = np.minimum(red, 0) * live[..., None], np.minimum(red_ampl, 0) * live[..., None]   # disregard (:248-252)
        score = 2 + R.sum(0)                                               # reset() returns counts = 2 as the initial score
        ep_len = 1 + live.sum(0)
        gotten = score - 2
        pct = gotten / ep_len * 100
        scores.append(score.copy())                                        # the shipped logs print raw scores as "avg score"
        agent_rew.append(pct)
        avg_score, med_score = np.mean(scores, 0), np.median(scores, 0)
        avg_r, med_r = np.mean(agent_rew[-100:], 0), np.median(agent_rew[-100:], 0)
        say("LOCAL ENV OUTPUTS: ")
        ep_metrics = []
        for i in range(n):
            L = int(min(max_len[i], T))
            a, b = red_d[:L, i], red_ampl_d[:L, i]
            nz = a[a != 0]
            m = dict(score=score[i], gotten_score=gotten[i], reward_percentage=pct[i],
                     torque_reduction_avg=a.mean(0), torque_reduction_median=np.median(a, 0), torque_reduction_std=a.std(0),
                     all_axes_pct=all_axes[:, i].sum() / (max_len[i] - 3) * 100, any_axis_pct=any_axes[:, i].sum() / (max_len[i] - 3) * 100,
                     ampl_reduction_avg=b.mean(0), ampl_reduction_median=np.median(b, 0), ampl_reduction_std=b.std(0),
                     ampl_occurrence_pct=ampl_neg[:, i].sum() / max_len[i] * 100,
                     suppression_avg=nz.mean() if nz.size else float("nan"), suppression_median=np.median(nz) if nz.size else float("nan"),
                     total_ampl_suppression=red_total_ep[:L, i].mean(), torque_places=seq.astype(np.float64), torque_max=torque_max[i])
            ep_metrics.append(m)
            say("TREMOR", i, "statistics: ")
            say(" score %.3f" % score[i], "avg score %.3f" % avg_score[i], "median score %.3f" % med_score[i],
                "max score %.3f" % np.max(np.array(scores)[:, i]), "std of scores %.3f" % np.std(np.array(scores)[-100:, i], axis=0))
            say("Reward achieved by the agent: %.3f, Score in percent to max: %.3f, avg rewards in percent: %.3f, median rewards in percent: %.3f"
                % (gotten[i], pct[i], avg_r[i], med_r[i]))
            say("Tremor torque reduction metrics:")
            say(" Tremor reduction ep avg", _fmt(m["torque_reduction_avg"]), "\n Tremor reduction ep median", _fmt(m["torque_reduction_median"]),
                "\n Tremor reduction ep std", _fmt(m["torque_reduction_std"]))
            say("Tremor reduction occurred in", m["all_axes_pct"], "% of all the generated tremor axis in all timesteps")
            say("Some sort of tremor reduction occurred along any axis in", m["any_axis_pct"], "% of the episode")
            say("Tremor amplitude reduction metrics:")
            say(" Tremor reduction ep avg", _fmt(m["ampl_reduction_avg"]), "\n Tremor reduction ep median", _fmt(m["ampl_reduction_median"]),
                "\n Tremor reduction ep std", _fmt(m["ampl_reduction_std"]))
            say("Tremor amplitude reduction occurred in", m["ampl_occurrence_pct"], "% of all timesteps")
            say("Overall tremor suppression avg in the angle axis: ", m["suppression_avg"], "%")
            say("Overall tremor suppression median in the angle axis: ", m["suppression_median"], "%")
            say("Total tremor amplitude suppression", m["total_ampl_suppression"], "%")
            say("For torque places: %s, With maximum Nm of: %s" % (_fmt(m["torque_places"]), _fmt(m["torque_max"])))
            say()
        metrics.append(ep_metrics)
    env.close()
    text = buf.getvalue()
    if out is not None:
        if hasattr(out, "write"):
            out.write(text)
        else:
            with open(out, "w") as f:
                f.write(text)
    return metrics, text
