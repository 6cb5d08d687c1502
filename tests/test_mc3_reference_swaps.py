"""CPU: `mc3.swap_chains` against the UNMODIFIED reference's `MCMCGenerative.swap_chains` (mcmc_generative.py:242-280)
under a shared decision stream (fixtures: oracle/make_golden_swaps.py, 40 scripted cases of 2-11 chains, 1-5 swap
rounds of 1-4 swaps): same partner indices and accept uniforms in, same `chain_idx`, `n_swaps`, `accepted_swaps` out."""
import os

import numpy as np

from conftest import GOLDEN
from contact_zones_b200.mc3 import swap_chains


class ScriptedRng:
    """Stands in for the numpy Generator of `swap_chains`: replays the recorded partner indices / uniforms."""

    def __init__(self, picks, unis):
        self.picks, self.unis = iter(picks), iter(unis)

    def integers(self, lo, hi):
        v = int(next(self.picks))
        assert lo <= v < hi
        return v

    def random(self):
        return float(next(self.unis))


def test_swap_chains_equals_reference():
    g = dict(np.load(os.path.join(GOLDEN, 'swaps.npz')))
    total = 0
    for i in range(int(g['n_cases'])):
        n, rounds, k = int(g[f'c{i}_n']), int(g[f'c{i}_rounds']), int(g[f'c{i}_chain_swaps'])
        chain_idx = list(range(n))
        stats = {'n_swaps': 0, 'accepted_swaps': 0}
        for r in range(rounds):
            rng = ScriptedRng(g[f'c{i}_picks'][r], g[f'c{i}_unis'][r])
            swap_chains(chain_idx, g[f'c{i}_lls'][r], g[f'c{i}_priors'][r], k, rng, stats)
            assert chain_idx == list(g[f'c{i}_idx_after'][r]), (i, r)
        assert stats['n_swaps'] == int(g[f'c{i}_n_swaps']) == rounds * k
        assert stats['accepted_swaps'] == int(g[f'c{i}_accepted'])
        total += stats['accepted_swaps']
    assert total > 50          # the fixtures exercise both outcomes
