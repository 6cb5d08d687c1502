"""TEST INFRASTRUCTURE ONLY -- golden vectors for the MC3 label swap (row a12): the UNMODIFIED reference's
`MCMCGenerative.swap_chains` (mcmc_generative.py:242-280) driven by a scripted decision stream.

The reference draws `np.random.choice(range(1, n_chains), 1)` for the partner and `random.random()` for the accept
test; both are replaced by a recorded stream here, so that `contact_zones_b200.mc3.swap_chains` can be fed the SAME
partner indices and uniforms (tests/test_mc3_reference_swaps.py) and must end with the same `chain_idx`, `n_swaps` and
`accepted_swaps`.   Run in the build container:   python oracle/make_golden_swaps.py
"""
import os
import random
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)

import refshim  # noqa: E402

ref = refshim.modules()
gen = ref.mcmc_generative


class _Stub(gen.MCMCGenerative):
    """swap_chains only needs chain_idx / n_chains / chain_swaps / statistics and likelihood() / prior()."""

    def __init__(self, n, chain_swaps):
        self.n_chains = n
        self.chain_idx = list(range(self.n_chains))
        self.chain_swaps = chain_swaps
        self.statistics = {'n_swaps': 0, 'accepted_swaps': 0}
        self._lls = self._priors = None

    def likelihood(self, sample, chain):
        return self._lls[chain]

    def prior(self, sample, chain):
        return self._priors[chain]

    # abstract in the reference class
    def generate_initial_sample(self, c=0):
        raise NotImplementedError

    def get_operators(self, operators):
        raise NotImplementedError


class _S:
    def everything_changed(self):
        pass


def main():
    rng = np.random.default_rng(2024)
    cases = []
    for case in range(40):
        n = int(rng.integers(2, 12))
        rounds = int(rng.integers(1, 6))
        chain_swaps = int(rng.integers(1, 5))
        spread = float(rng.choice([0.5, 3.0, 30.0]))
        picks = rng.integers(1, n, size=(rounds, chain_swaps))
        unis = rng.random((rounds, chain_swaps))
        lls = rng.normal(-1000.0, spread, size=(rounds, n))
        priors = rng.normal(-20.0, spread / 3, size=(rounds, n))
        stub = _Stub(n, chain_swaps)
        script = {'p': iter(()), 'u': iter(())}
        orig_choice, orig_random = np.random.choice, gen._random.random
        np.random.choice = lambda a, size=None, **kw: np.array([next(script['p'])])
        gen._random.random = lambda: next(script['u'])
        idx_after = []
        try:
            for r in range(rounds):
                stub._lls, stub._priors = lls[r], priors[r]
                script['p'], script['u'] = iter(picks[r]), iter(unis[r])
                stub.swap_chains([_S() for _ in range(n)])
                idx_after.append(list(stub.chain_idx))
        finally:
            np.random.choice, gen._random.random = orig_choice, orig_random
        cases.append(dict(n=n, rounds=rounds, chain_swaps=chain_swaps, picks=picks, unis=unis, lls=lls, priors=priors,
                          idx_after=np.array(idx_after), n_swaps=stub.statistics['n_swaps'],
                          accepted=stub.statistics['accepted_swaps']))
    save = {'n_cases': np.int64(len(cases))}
    for i, c in enumerate(cases):
        for k, v in c.items():
            save[f'c{i}_{k}'] = np.asarray(v)
    out = os.path.join(ROOT, 'tests', 'golden', 'swaps.npz')
    np.savez_compressed(out, **save)
    print('wrote', out, os.path.getsize(out), 'B;', sum(int(c['accepted']) for c in cases), 'accepted of',
          sum(int(c['n_swaps']) for c in cases))


if __name__ == '__main__':
    main()
