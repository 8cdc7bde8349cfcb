"""CPU restatement of the policy-input stage that sits on the other side of ``obs_buf``.

TEST INFRASTRUCTURE ONLY (see oracle/dexstep_oracle.py): the checker for ``DexOutBuffers.obs_norm``.

PARITY UNPINNED.  The algorithm lives in a third-party dependency that is absent from /root/reference and from this
image: rl_games (setup.py:23 pins ``rl-games>=1.6.0``), module ``rl_games/algos_torch/running_mean_std.py``, class
``RunningMeanStd`` - the reference only switches it on (cfg/train/BaseTaskPPO.yaml:44 ``normalize_input: True``) and
reloads its buffers on checkpoint hot-reload (rl/rl_games_patches.py:143-150).  What follows restates the published
1.6.x algorithm for a flat observation (``insize=(obs_dim,)``, ``per_channel=False``, ``norm_only=False``,
``epsilon=1e-05``): float64 buffers running_mean (zeros), running_var (ones), count (ones);

    training:  mean, var = x.mean(0), x.var(0)            (unbiased variance over the batch of envs)
               running_* <- parallel-variance merge of (running_mean, running_var, count) with (mean, var, N)
    always:    y = (x - running_mean.float()) / sqrt(running_var.float() + epsilon);  y = clamp(y, -5, 5)

No golden vector of the reference covers it (the reference has no test that runs rl_games), hence "unpinned".
"""
from __future__ import annotations

import torch


class RunningMeanStd:
    def __init__(self, obs_dim: int, epsilon: float = 1e-5, clip: float = 5.0):
        self.epsilon, self.clip = epsilon, clip
        self.running_mean = torch.zeros(obs_dim, dtype=torch.float64)
        self.running_var = torch.ones(obs_dim, dtype=torch.float64)
        self.count = torch.ones((), dtype=torch.float64)
        self.training = False

    def update(self, x: torch.Tensor):
        """_update_mean_var_count_from_moments: Chan et al. merge of the batch moments into the running ones."""
        batch_mean, batch_var, batch_count = x.mean(0), x.var(0), x.shape[0]
        delta = batch_mean - self.running_mean
        tot = self.count + batch_count
        new_mean = self.running_mean + delta * batch_count / tot
        m2 = self.running_var * self.count + batch_var * batch_count + delta ** 2 * self.count * batch_count / tot
        self.running_mean, self.running_var, self.count = new_mean, m2 / tot, tot

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.training:
            self.update(x)
        y = (x - self.running_mean.float()) / torch.sqrt(self.running_var.float() + self.epsilon)
        return torch.clamp(y, min=-self.clip, max=self.clip)
