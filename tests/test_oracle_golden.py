"""CPU: the numpy oracle against the golden vectors generated from the imported reference
(oracle/make_golden.py).  Tolerances: floats 1e-12 relative (same arithmetic, different summation order at
most), integer count tables exact."""
import numpy as np
import pytest

import sbayes_oracle as orc
from conftest import golden_cases
from helpers import load_case, rel_err


@pytest.mark.parametrize('name', golden_cases())
def test_likelihood(name):
    c = load_case(name)
    g = c['g']
    args = (c['features'], c['families'], c['inheritance'], g['zones'], g['weights'], g['p_global'], g['p_zones'], c['p_families'])
    assert rel_err(orc.log_likelihood(*args, source=None), g['ref_ll_nosrc']) < 1e-12
    assert rel_err(orc.log_likelihood(*args, source=c['source']), g['ref_ll_src']) < 1e-12
    assert np.max(np.abs(orc.source_posterior(*args) - g['ref_post'])) < 1e-14
    w = orc.normalized_weights(g['weights'], orc.has_components(g['zones'], c['families'], c['inheritance']))
    assert np.max(np.abs(w - g['ref_wnorm'])) < 1e-15


@pytest.mark.parametrize('name', golden_cases())
def test_prior(name):
    c = load_case(name)
    g = c['g']
    p = orc.log_prior(g['zones'], g['p_global'], c['p_families'], g['states'], c['inheritance'], c['prior_cfg'],
                      cost_mat=c['cost'], alpha_global=g.get('alpha_global'), alpha_families=g.get('alpha_fam'))
    assert rel_err(p, g['ref_prior']) < 1e-12 or abs(p - g['ref_prior']) < 1e-12
    assert abs(orc.size_prior(g['zones'], c['size_type']) - g['ref_size']) < 1e-10
    if c['geo_type'] == 'cost_based':
        assert rel_err(orc.geo_prior_cost_based(g['zones'], c['cost'], c['geo_scale']), g['ref_geo']) < 1e-12


@pytest.mark.parametrize('name', golden_cases())
def test_counts_exact(name):
    c = load_case(name)
    g = c['g']
    src = c['source']
    assert np.array_equal(orc.source_counts(src, g['zones'].any(0)), g['ref_cnt_src_zone'])
    assert np.array_equal(orc.state_counts_global(c['features'], src), g['ref_cnt_glob'])
    for z in range(c['Z']):
        assert np.array_equal(orc.state_counts_zone(c['features'], src, g['zones'][z]), g['ref_cnt_zone'][z])
    if c['inheritance']:
        assert np.array_equal(orc.source_counts(src, c['families'].any(0)), g['ref_cnt_src_fam'])
        for k in range(c['families'].shape[0]):
            assert np.array_equal(orc.state_counts_family(c['features'], src, c['families'][k]), g['ref_cnt_fam'][k])


@pytest.mark.parametrize('name', golden_cases())
def test_move_probabilities(name):
    c = load_case(name)
    g = c['g']
    assert len(g['ref_moves']) > 0
    for kind, z, sn, sr, lq, lqb in g['ref_moves']:
        kind, z, sn, sr = int(kind), int(z), int(sn), int(sr)
        if kind == 0:
            o = orc.grow_log_q(g['zones'], z, sn, c['adj'], 0.85)
        elif kind == 1:
            o = orc.shrink_log_q(g['zones'], z, sr, c['adj'], 0.85)
        else:
            o = orc.swap_log_q(g['zones'], z, sn, sr, c['adj'], 0.85)
        assert abs(o[0] - lq) < 1e-12 and abs(o[1] - lqb) < 1e-12
