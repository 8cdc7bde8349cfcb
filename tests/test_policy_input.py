"""CPU tests of the policy-input oracle (rl_games RunningMeanStd restatement, oracle/policy_input.py): the moment merge
is order-independent and equals the moments of the pooled data; eval mode leaves the statistics alone; the output is
clipped.  (Parity of this stage is unpinned: rl_games is not in /root/reference, see the module header.)"""
import torch

from oracle.policy_input import RunningMeanStd


def _pooled(prior_mean, prior_var, prior_count, batches):
    """Moments of {prior pseudo-sample} U batches computed directly in float64 (population-style merge of sums)."""
    n = prior_count
    s1 = prior_mean * prior_count
    s2 = (prior_var + prior_mean ** 2) * prior_count
    for b in batches:
        b = b.double()
        m, v, k = b.mean(0), b.var(0), b.shape[0]
        n = n + k
        s1 = s1 + m * k
        s2 = s2 + (v + m ** 2) * k
    mean = s1 / n
    return mean, s2 / n - mean ** 2, n


def test_update_equals_pooled_moments():
    g = torch.Generator().manual_seed(0)
    d = 37
    b1 = torch.randn(1000, d, generator=g) * 3.0 + 1.5
    b2 = torch.randn(64, d, generator=g) * 0.1 - 4.0
    b3 = torch.rand(4096, d, generator=g)
    r = RunningMeanStd(d)
    r.training = True
    for b in (b1, b2, b3):
        r.forward(b)
    mean, var, n = _pooled(torch.zeros(d, dtype=torch.float64), torch.ones(d, dtype=torch.float64), 1.0, (b1, b2, b3))
    assert float(r.count) == n == 1.0 + 1000 + 64 + 4096
    assert torch.allclose(r.running_mean, mean, rtol=1e-5, atol=1e-6)  # batch moments are fp32 like rl_games'
    assert torch.allclose(r.running_var, var, rtol=1e-5, atol=1e-6)
    r2 = RunningMeanStd(d)
    r2.training = True
    for b in (b3, b1, b2):
        r2.forward(b)
    assert torch.allclose(r2.running_mean, r.running_mean, rtol=1e-9, atol=1e-9)
    assert torch.allclose(r2.running_var, r.running_var, rtol=1e-9, atol=1e-9)


def test_eval_forward_is_frozen_and_clipped():
    d = 5
    r = RunningMeanStd(d)
    r.running_mean = torch.tensor([0.0, 1.0, -2.0, 0.5, 0.0], dtype=torch.float64)
    r.running_var = torch.tensor([1.0, 4.0, 1e-8, 0.25, 1.0], dtype=torch.float64)
    x = torch.tensor([[0.0, 3.0, -2.0, 0.0, 100.0], [1.0, 1.0, -1.9, 1.0, -100.0]])
    before = (r.running_mean.clone(), r.running_var.clone(), r.count.clone())
    y = r.forward(x)
    assert all(torch.equal(a, b) for a, b in zip(before, (r.running_mean, r.running_var, r.count)))
    den = torch.sqrt(r.running_var.float() + 1e-5)
    assert torch.equal(y, torch.clamp((x - r.running_mean.float()) / den, -5.0, 5.0))
    assert y[0, 4] == 5.0 and y[1, 4] == -5.0 and y[1, 2] == 5.0 and y[0, 1] == (3.0 - 1.0) / den[1]
