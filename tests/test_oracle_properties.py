"""Property tests of the CPU oracle itself (tests/oracle.py) — the spec every transport and kernel is compared
with.  They are the reference's two unit-test properties (test/test_AllReduceSGD.lua:37-39: exact equality after
synchronizeParameters; test/test_AllReduceEA.lua:36-39: nodes converge to within 1e-6) plus the conservation law
of the asynchronous exchange, over random node counts and uneven step counts."""
import torch
from hypothesis import given, settings, strategies as st

import oracle


@settings(max_examples=40, deadline=None)
@given(st.sampled_from([2, 4, 8]), st.integers(0, 2**31 - 1), st.integers(1, 4))
def test_sgd_nodes_end_bit_identical_whatever_the_step_counts(n, seed, epochs):
    g = torch.Generator().manual_seed(seed)
    sim = oracle.SgdNodeSim([torch.randn(7, generator=g) for _ in range(n)])
    sim.synchronize([0] * n)
    for _ in range(epochs):
        steps = torch.randint(4, 14, (n,), generator=g).tolist()          # 4..13 like the reference test
        sim.epoch(steps, lambda r, s, k: torch.full((7,), 1.0 / k))
        for r in range(1, n):
            assert torch.equal(sim.params[0], sim.params[r])
        assert torch.isfinite(sim.params[0]).all()


@settings(max_examples=25, deadline=None)
@given(st.sampled_from([2, 4, 8]), st.integers(0, 2**31 - 1))
def test_ea_nodes_contract_onto_the_center(n, seed):
    """tau = 3, a random walk whose steps halve.  Differences between nodes shrink by (1 - alpha) per round and the
    distance to the center by |1 - alpha (1 + n)|; alpha = 1.6 / (1 + n) keeps both modes contractive for every n
    (the reference's alpha = 0.4 is only contractive in the first mode for n = 8), so everything collapses."""
    g = torch.Generator().manual_seed(seed)
    sim = oracle.EaNodeSim([torch.randn(7, generator=g) for _ in range(n)], tau=3, alpha=1.6 / (1 + n))
    sim.synchronize_parameters()
    slow = [1.0] * n

    def move(r, _i):
        v = torch.randn(7, generator=g) / slow[r]
        slow[r] *= 2
        return v

    for _ in range(5):
        steps = torch.randint(45, 54, (n,), generator=g).tolist()
        sim.epoch(steps, move)
    for r in range(n):
        assert float((sim.xs[r] - sim.xs[0]).abs().max()) < 1e-4
    spread = max(float((x - sim.center).abs().max()) for x in sim.xs)
    assert spread < 1e-3


@settings(max_examples=50, deadline=None)
@given(st.integers(1, 7), st.floats(0.01, 0.9), st.integers(0, 2**31 - 1))
def test_async_exchange_conserves_mass(clients, alpha, seed):
    g = torch.Generator().manual_seed(seed)
    center = torch.randn(5, generator=g, dtype=torch.float64)
    xs = [torch.randn(5, generator=g, dtype=torch.float64) for _ in range(clients)]
    total = center + sum(xs)
    order = torch.randint(0, clients, (30,), generator=g).tolist()
    for c in order:
        xs[c], center = oracle.async_client_sync(xs[c], center, alpha)
    torch.testing.assert_close(center + sum(xs), total, rtol=1e-12, atol=1e-12)


def test_winner_rule():
    assert oracle.winner([0, 0, 0]) == 0                 # nobody stepped: node 1 scatters (lua/AllReduceSGD.lua:50-52)
    assert oracle.winner([3, 9, 9, 2]) == 2              # ties: last index of the ascending sort (:41-43)
    assert oracle.winner([5, 1]) == 0
