"""Per-env, per-episode suppression statistics of the training script - means, MEDIANS and standard deviations over the
episode's time steps (Simulation/Exoskeleton_agent_train.py:150-205 per step, :219-265 per episode) - kept on the device.

The step kernel already reduces the script's sums and occurrence counters into exo_stats; medians and
standard deviations need the episode's history, so this keeps [T, N, 7] tensors of the per-step percentages (the script's
`tremor_reduction_full_ep` / `tremor_ampl_reduction_full_ep` / `tremor_ampl_total_reduction_full_ep` lists) next to the env
and reduces them with torch at episode end. Memory is 15 floats per env-step: meant for the script's population sizes
(8 ... a few thousand envs), not for the 65 536-env benchmark."""
import torch


class EpisodeSuppressionLog(object):
    def __init__(self, n_envs, max_steps, device, disregard=True):
        self.n, self.T, self.disregard = int(n_envs), int(max_steps), bool(disregard)
        dev = torch.device(device)
        self.red = torch.zeros(self.T, self.n, 7, device=dev, dtype=torch.float64)        # tremor_reduction_full_ep
        self.red_ampl = torch.zeros(self.T, self.n, 7, device=dev, dtype=torch.float64)   # tremor_ampl_reduction_full_ep
        self.ampl_total = torch.zeros(self.T, self.n, device=dev, dtype=torch.float64)    # tremor_ampl_total_reduction_full_ep
        self.when = torch.zeros(self.n, 2, device=dev, dtype=torch.float64)               # tremor_when_reduction
        self.when_ampl = torch.zeros(self.n, 2, device=dev, dtype=torch.float64)          # tremor_when_ampl_reduction
        self.t = 0

    def reset(self):
        for x in (self.red, self.red_ampl, self.ampl_total, self.when, self.when_ampl):
            x.zero_()
        self.t = 0

    @staticmethod
    def _pct(a, b):
        """(|a| - |b|) / |b| * 100 with nan / +-inf -> 0 (train:150-155)."""
        r = (a.abs() - b.abs()) / b.abs() * 100.0
        return torch.nan_to_num(r, nan=0.0, posinf=0.0, neginf=0.0)

    def add(self, info, live):
        """One control step. info: the env's info tensors ([N,7] torque_val, tremor_torque_val, ampl_val, tremor_ampl_val and
        [N] amp_metric = the script's tremor_reduction_ampl_total); live: bool [N], envs that stepped (train:139)."""
        if self.t >= self.T:
            raise IndexError("episode longer than the log")
        live = live.to(torch.bool)
        red = self._pct(info["torque_val"].double(), info["tremor_torque_val"].double())
        red_a = self._pct(info["ampl_val"].double(), info["tremor_ampl_val"].double())
        tot = info["amp_metric"].double()
        lv = live.double()
        self.when[:, 0] += (red[:, :4] >= 0).sum(1) * lv                 # train:172-173
        self.when[:, 1] += (red[:, :4] < 0).sum(1) * lv
        self.when_ampl[:, 1] += (tot < 0) * lv                           # train:179-184
        self.when_ampl[:, 0] += (~(tot < 0)) * lv
        tot_ep = torch.where(tot < 0, tot, torch.zeros_like(tot))        # only negative totals are stored (:181)
        if self.disregard:                                               # train:187-191
            red = torch.where(red > 0, torch.zeros_like(red), red)
            red_a = torch.where(red_a > 0, torch.zeros_like(red_a), red_a)
        m = live[:, None]
        self.red[self.t] = torch.where(m, red, torch.zeros_like(red))    # finished envs log zeros (:198-200)
        self.red_ampl[self.t] = torch.where(m, red_a, torch.zeros_like(red_a))
        self.ampl_total[self.t] = torch.where(live, tot_ep, torch.zeros_like(tot_ep))
        self.t += 1

    def summary(self, max_episode_lengths):
        """The script's per-env console figures (train:239-265) as tensors; env i's statistics run over its first
        max_episode_lengths[i] logged steps, as the script slices them."""
        L = torch.as_tensor(max_episode_lengths, device=self.red.device).clamp(max=self.t)
        out = {k: [] for k in ("tremor_avg", "tremor_median", "tremor_std", "tremor_ampl_avg", "tremor_ampl_median", "tremor_ampl_std",
                               "tremor_sup_avg", "tremor_sup_median", "tremor_ampl_sup_avg", "tremor_ampl_total_avg")}
        for i in range(self.n):
            a, b, c = self.red[:int(L[i]), i], self.red_ampl[:int(L[i]), i], self.ampl_total[:int(L[i]), i]
            out["tremor_avg"].append(a.mean(0)); out["tremor_std"].append(a.std(0, unbiased=False))
            out["tremor_median"].append(_np_median(a))
            out["tremor_ampl_avg"].append(b.mean(0)); out["tremor_ampl_std"].append(b.std(0, unbiased=False))
            out["tremor_ampl_median"].append(_np_median(b))
            nz, nzb = a[a != 0], b[b != 0]
            out["tremor_sup_avg"].append(nz.mean() if nz.numel() else a.new_tensor(float("nan")))
            out["tremor_sup_median"].append(_np_median(nz[:, None])[0] if nz.numel() else a.new_tensor(float("nan")))
            out["tremor_ampl_sup_avg"].append(nzb.mean() if nzb.numel() else a.new_tensor(float("nan")))
            out["tremor_ampl_total_avg"].append(c.mean())
        res = {k: torch.stack(v) for k, v in out.items()}
        res["tremor_occurrence_pct"] = self.when[:, 1] / self.when.sum(1) * 100.0
        res["tremor_ampl_occurrence_pct"] = self.when_ampl[:, 1] / self.when_ampl.sum(1) * 100.0
        return res


def _np_median(x):
    """np.median over axis 0 (mean of the two middle values for an even count; torch.median returns the lower one)."""
    n = x.shape[0]
    s = torch.sort(x, dim=0).values
    return (s[(n - 1) // 2] + s[n // 2]) / 2.0
