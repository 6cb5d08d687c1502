"""CPU: host-side initial state of the sampler (a18, zone_sampling.py:935-1233) and the operator table handling --
no device needed (the context is opened lazily by generate_samples)."""
import numpy as np
import pytest
from scipy.sparse.csgraph import connected_components

from helpers import PRECISION, case_data, case_model_config, load_case


def _sampler(c, **kw):
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import ZoneMCMC
    data = case_data(c)
    model = Model(data, case_model_config(c))
    ops = {'grow_zone': 0.5, 'shrink_zone': 0.5, 'gibbsish_sample_zones': 0.0}
    return ZoneMCMC(data=data, model=model, operators=ops, var_proposal=PRECISION, p_grow_connected=0.85,
                    initial_size=5, n_chains=3, seed=3, **kw), data, model


def test_initial_sample_host():
    c = load_case('C_samerica_like')
    smp, data, model = _sampler(c)
    s = smp._initial_sample_host(0)
    assert s.zones.shape == (c['Z'], c['N']) and (s.zones.sum(axis=1) == 5).all() and (s.zones.sum(axis=0) <= 1).all()
    for z in s.zones:        # grown through neighbours => connected in the adjacency graph
        idx = np.flatnonzero(z)
        assert connected_components(c['adj'][idx][:, idx], directed=False)[0] == 1
    assert np.allclose(s.weights, 1. / c['C'])
    app = data.states.astype(bool)
    for p in (s.p_global[0], s.p_zones[0], s.p_families[0]):
        assert np.allclose(p.sum(axis=-1), 1.) and (p[~app] == 0).all() and (p[app] > 0).all()
    cnt = data.features.sum(axis=0).astype(float)
    cnt[app] += 1
    assert np.allclose(s.p_global[0], cnt / cnt.sum(axis=1, keepdims=True))
    # same seed, same chain -> same initial state, whatever was generated before (sharding independence)
    smp._initial_sample_host(2)
    assert np.array_equal(smp._initial_sample_host(0).zones, s.zones)
    assert not np.array_equal(smp._initial_sample_host(1).zones, s.zones)


def test_initial_sample_continues_from_previous():
    from contact_zones_b200.sampling import Sample
    c = load_case('C_samerica_like')
    g = c['g']
    prev = Sample(g['zones'][:1], g['weights'], g['p_global'], g['p_zones'][:1], c['p_families'])
    smp, _, _ = _sampler(c, initial_sample=prev)
    s = smp._initial_sample_host(0)
    assert np.array_equal(s.zones[0], g['zones'][0]) and np.array_equal(s.p_zones[0], g['p_zones'][0])
    assert (s.zones[1:].sum(axis=1) == 5).all() and (s.zones.sum(axis=0) <= 1).all()
    assert np.allclose(s.weights, g['weights'] / g['weights'].sum(-1, keepdims=True))
    assert np.array_equal(s.p_families, c['p_families'])


def test_unknown_operator_is_an_error():
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import ZoneMCMC
    c = load_case('C_samerica_like')
    data = case_data(c)
    model = Model(data, case_model_config(c))
    with pytest.raises(AttributeError):
        ZoneMCMC(data=data, model=model, operators={'teleport_zone': 0.5}, var_proposal=PRECISION,
                 p_grow_connected=0.85, initial_size=5, n_chains=1)
    # sample_from_prior is part of the constructor contract (mcmc_generative.py:33-36, 115-116): likelihood() is 0
    smp = ZoneMCMC(data=data, model=model, operators={'grow_zone': 1.}, var_proposal=PRECISION, p_grow_connected=0.85,
                   initial_size=5, n_chains=1, sample_from_prior=True)
    assert smp.sample_from_prior and smp.likelihood(None, 0) == 0.
