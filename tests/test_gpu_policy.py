"""
Fused policy forward (csrc/swarm_policy.cu) against the oracle's fp32 restatement of the
reference's action path (networks/flax_network.py:153-195, sampling_strategies/gumbel_distribution.py:19-42,
network of CI/espresso_tests/integration_tests/test_rl_trainers.py:17-26).  flax / JAX cannot be imported
offline, so the deterministic outputs are pinned to the textbook Dense / softmax arithmetic (tolerance 2e-5,
fp32 summation order) and the sampler to its defining property: action frequencies follow softmax(logits).
"""
import numpy as np
import pytest

from helpers import jittered_lattice, species_for, wca_for

pytestmark = pytest.mark.gpu
TOL = 2e-5


def _swarm(n, types=None, seed=42):
    from swarmrl_b200.device import DeviceSwarm

    rng = np.random.default_rng(1)
    sp, wca = species_for([1.0, 1.0]), wca_for([1.0, 1.0])
    d = DeviceSwarm(2, (500.0,) * 3, True, 1e-3, 0.0, seed, sp, wca)
    d.set_state(jittered_lattice(n, 500.0, 2, rng), np.zeros(n, dtype=np.int32) if types is None else types, theta=np.zeros(n))
    return d


def _params(rng, D, H=128, A=4, scale=0.3):
    return dict(w1=rng.normal(size=(D, H)) * scale, b1=rng.normal(size=H) * 0.1, wa=rng.normal(size=(H, A)) * scale,
                ba=rng.normal(size=A) * 0.1, wv=rng.normal(size=H) * scale, bv=rng.normal(size=1))


@pytest.mark.parametrize("D,A,H", [(1, 4, 128), (5, 4, 128), (12, 7, 64), (32, 16, 200)])
def test_logits_value_logprob_and_sampling_match_oracle(D, A, H):
    import torch

    from oracle.oracle import policy_forward
    from swarmrl_b200.agents import FusedActorCritic

    rng = np.random.default_rng(D * 100 + A)
    n = 777
    d = _swarm(n)
    P = _params(rng, D, H, A)
    obs = rng.normal(size=(n, D)) * 2.0
    obs_dev = torch.from_numpy(obs).to(d.dev)
    pol = FusedActorCritic(P, d)
    for call, sample in ((0, False), (0, True), (5, True)):
        idx, logp, val, lg = d.policy_forward(pol._dev, obs_dev, None, n, call, sample, want_logits=True)
        o_idx, o_logp, o_val, o_lg = policy_forward(obs, P, seed=42, call=call, sample=sample)
        assert np.allclose(lg.cpu().numpy(), o_lg, rtol=TOL, atol=TOL * 10)
        assert np.allclose(val.cpu().numpy(), o_val, rtol=TOL, atol=TOL * 10)
        got = idx.cpu().numpy()
        same = got == o_idx
        assert same.mean() > 0.995          # a near-tie of two perturbed logits may flip under fp32 reordering
        assert np.allclose(logp.cpu().numpy()[same], o_logp[same], rtol=1e-4, atol=1e-4)


def test_gumbel_max_frequencies_follow_softmax():
    import torch

    from swarmrl_b200.agents import FusedActorCritic

    n, D, A = 50000, 3, 4
    d = _swarm(n)
    rng = np.random.default_rng(3)
    P = _params(rng, D, 32, A)
    obs = np.tile(rng.normal(size=(1, D)), (n, 1))          # every agent sees the same observable
    pol = FusedActorCritic(P, d)
    idx, logp, val, lg = d.policy_forward(pol._dev, torch.from_numpy(obs).to(d.dev), None, n, 11, True, want_logits=True)
    l0 = lg[0].double().cpu().numpy()
    p = np.exp(l0 - l0.max()); p /= p.sum()
    freq = np.bincount(idx.cpu().numpy(), minlength=A) / n
    assert np.all(np.abs(freq - p) < 4.0 * np.sqrt(p * (1 - p) / n) + 1e-3), (freq, p)
    # different calls give different draws, the same call the same draws
    i2, _, _ = d.policy_forward(pol._dev, torch.from_numpy(obs).to(d.dev), None, n, 12, True)
    i3, _, _ = d.policy_forward(pol._dev, torch.from_numpy(obs).to(d.dev), None, n, 11, True)
    assert (i2 != idx).any() and torch.equal(i3, idx)


def test_chosen_action_is_written_for_the_agent_type_only():
    """manage_forces fused in (espresso.py:1227-1235): f_swim / torque of the selected type get the table rows,
    the other type is left alone; flax-style parameter tree accepted."""
    import torch

    from swarmrl_b200.agents import FusedActorCritic

    n = 1000
    rng = np.random.default_rng(9)
    types = (rng.uniform(size=n) < 0.4).astype(np.int32)
    d = _swarm(n, types)
    P = _params(rng, 5)
    tree = {"Dense_0": {"kernel": P["w1"], "bias": P["b1"]}, "Dense_1": {"kernel": P["wv"].reshape(-1, 1), "bias": P["bv"]},
            "Dense_2": {"kernel": P["wa"], "bias": P["ba"]}}
    pol = FusedActorCritic(tree, d, sample=False)
    row_of, rows = d.row_map(1)
    obs = torch.from_numpy(rng.normal(size=(rows, 5))).to(d.dev)
    f_tab = torch.tensor([10.0, 0.0, 0.0, 0.0], dtype=torch.float64, device=d.dev)
    t_tab = torch.tensor([[0, 0, 0], [0, 0, 7.0], [0, 0, -7.0], [0, 0, 0]], dtype=torch.float64, device=d.dev)
    d.f_swim[:] = -1.0
    idx, logp = pol.act(d, obs, row_of, rows, f_tab, t_tab)
    f = d.f_swim[:n].cpu().numpy(); tz = d.torque[2, :n].cpu().numpy(); k = idx.cpu().numpy()
    sel = types == 1
    assert np.all(f[~sel] == -1.0)
    assert np.array_equal(f[sel], f_tab.cpu().numpy()[k]) and np.array_equal(tz[sel], t_tab.cpu().numpy()[k, 2])
    assert pol.call == 1 and pol.last_value.shape[0] == rows
