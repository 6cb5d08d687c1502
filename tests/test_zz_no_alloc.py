"""GPU: `sb_eval_likelihood_noalloc` (scratch in a file-scope `static __device__` array, no cudaMalloc in the call)
against the existing entry points `sb_eval_likelihood_async` + `sb_read_ll`, through the C ABI, on two calls in a row
that use different chain ranges -- the shared array keeps the first call's partial sums and tickets.

Collected last on purpose (file name)."""
import numpy as np
import pytest

from contact_zones_b200 import synthetic
from contact_zones_b200._lib import SbError
from contact_zones_b200.engine import Engine

pytestmark = pytest.mark.gpu
RTOL = 1e-9          # the likelihood bar of tests/test_gpu_parity.py


def _engine(n_sites, n_features, n_states, n_zones, n_fam, n_chains, seed):
    data = synthetic.make_data(n_sites, n_features, n_states, n_families=n_fam, seed=seed, with_dist=False)
    inh = n_fam > 0
    eng = Engine(data.features, data.states, data.families if inh else None, data.network['adj_mat'], None,
                 n_zones=n_zones, n_chains=n_chains, inheritance=inh, sample_source=False)
    for c in range(n_chains):
        st = synthetic.random_state(data, n_zones, 12, inh, False, seed=100 * seed + c)
        eng.set_state(c, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'] if inh else None)
    return eng


@pytest.mark.parametrize('shape', [(3000, 60, 4, 3, 0, 6), (700, 37, 7, 2, 5, 9)])
def test_noalloc_matches_existing_entry_point_twice(shape, built_lib):
    eng = _engine(*shape, seed=7)
    n = shape[-1]
    eng.eval_likelihood_async(0, n)
    ref_all = eng.read_ll(0, n)
    assert np.all(np.isfinite(ref_all))
    # call 1: every chain; call 2, right behind it: a sub-range (another grid, so the stale scratch is laid out differently)
    got_all = eng.eval_likelihood_noalloc(0, n)
    got_sub = eng.eval_likelihood_noalloc(1, n - 2)
    eng.eval_likelihood_async(1, n - 2)
    ref_sub = eng.read_ll(1, n - 2)
    assert np.max(np.abs(got_all - ref_all) / np.abs(ref_all)) <= RTOL
    assert np.max(np.abs(got_sub - ref_sub) / np.abs(ref_sub)) <= RTOL
    assert np.max(np.abs(ref_sub - ref_all[1:n - 1]) / np.abs(ref_sub)) <= RTOL
    # a third call after the state of one chain changed
    eng.copy_state(0, n - 1)
    got3 = eng.eval_likelihood_noalloc(0, n)
    assert abs(got3[0] - ref_all[n - 1]) <= RTOL * abs(ref_all[n - 1])
    assert np.max(np.abs(got3[1:] - ref_all[1:]) / np.abs(ref_all[1:])) <= RTOL
    eng.close()


def test_noalloc_device_output_and_limits(built_lib):
    import torch
    eng = _engine(1200, 24, 4, 2, 0, 5, seed=3)
    eng.eval_likelihood_async(0, 5)
    ref = eng.read_ll(0, 5)
    out = torch.zeros(5, dtype=torch.float64, device='cuda')
    torch.cuda.synchronize()
    eng.eval_likelihood_noalloc(0, 5, device_ptr=out.data_ptr())      # stream-ordered on the context's stream
    eng.sync()
    assert np.max(np.abs(out.cpu().numpy() - ref) / np.abs(ref)) <= RTOL
    with pytest.raises(SbError):
        eng.eval_likelihood_noalloc(3, 5)                              # chain range out of bounds
    eng.close()
    # beyond the static scratch: more than 4096 chains in one call is refused with a clear message
    big = Engine(np.zeros((64, 4), np.uint8), np.ones((4, 2), bool), n_zones=0, n_chains=4100)
    with pytest.raises(SbError, match='4096'):
        big.eval_likelihood_noalloc(0, 4100)
    big.close()
