"""CPU restatement of the reference's RewardComponentObserver accumulation (dexhand_env/rl/reward_observer.py:160-293).

TEST INFRASTRUCTURE ONLY - the checker for the device-side accumulator that dex_post_physics fuses into the step
(DexOutBuffers.episode_sums / episode_lengths / stat_totals).  Pinned against the UNMODIFIED reference
observer by tests/test_observer_oracle.py (oracle/ref_harness.py runs it behind stubs of the two rl_games names it
imports).

Call order per control step, as the reference's docstring states it ("called after each training step"):
    after_steps()                       :160-174  episode_lengths += 1 ; sums[name] += extras["reward_components"][name]
    process_infos(infos, done_indices)  :137-158 -> _process_done_episodes_vectorized :190-239 -> _update_statistics :241-293
What is compared with the device accumulator are the reference's *cumulative* quantities:
    total_episodes / episodes_by_type[t]                      <-> stat_totals[cls][0]
    cumulative_sums[t][name]["steps"] (sum of done lengths)   <-> stat_totals[cls][1]
    cumulative_sums[t][name]["rewards"] (sum of done sums)    <-> stat_totals[cls][2 + row(name)]
(the reference skips the component "total", :254; the device keeps it as one more row).
"""
from __future__ import annotations

from typing import Dict, List

import torch

CLASSES = ("all", "success", "failure", "timeout")


class OracleObserver:
    def __init__(self, num_envs: int, component_names: List[str]):
        self.names = list(component_names)
        self.n = num_envs
        self.episode_lengths = torch.zeros(num_envs, dtype=torch.long)           # :107
        self.episode_sums = {k: torch.zeros(num_envs) for k in self.names}       # :120-122
        # [class][0] episodes, [class][1] sum of lengths, [class][2 + row] sum of the episode sums (float64 accumulators,
        # like the reference's Python floats)
        self.totals = torch.zeros(4, 2 + len(self.names), dtype=torch.float64)

    def after_steps(self, reward_components: Dict[str, torch.Tensor]):
        self.episode_lengths += 1                                                 # :167
        for k in self.names:                                                      # :176-188
            self.episode_sums[k] += reward_components[k].detach().cpu().float()

    def process_infos(self, infos: Dict[str, torch.Tensor], done_indices: torch.Tensor):
        if len(done_indices) == 0:                                                # :146
            return
        done = done_indices.cpu().long()
        masks = [None, infos["success"].cpu().bool(), infos["failure"].cpu().bool(), infos["timeout"].cpu().bool()]  # :196-206
        lengths = self.episode_lengths[done].float()                              # :209
        for c, m in enumerate(masks):
            sel = torch.ones(done.numel(), dtype=torch.bool) if m is None else m[done]   # :272-277
            self.totals[c, 0] += float(sel.sum())
            self.totals[c, 1] += float(lengths[sel].sum())
            for r, k in enumerate(self.names):
                self.totals[c, 2 + r] += float(self.episode_sums[k][done][sel].sum().item())   # :258-262, :281-292
        self.episode_lengths[done] = 0                                            # :217-219
        for k in self.names:
            self.episode_sums[k][done] = 0
