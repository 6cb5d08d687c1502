"""GPU: the reference-shaped objects (`contact_zones_b200.model`) through the C ABI.

`test_family_area_overlap` is the reference's own unit test (test/test_model.py:52-92) with our `Likelihood` /
`Sample` substituted -- same stub `Data` namedtuple, same assertions, seeded.  The golden tests check `Model`,
`Likelihood`, `Prior` and every prior term against values produced by the unmodified reference
(oracle/make_golden.py).  Bar: 1e-9 relative."""
from collections import namedtuple

import numpy as np
import pytest

from conftest import golden_cases
from helpers import case_data, case_model_config, load_case

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def close(a, b):
    return abs(a - b) <= RTOL * max(1.0, abs(b))


def test_family_area_overlap(built_lib):
    from contact_zones_b200.model import Likelihood
    from contact_zones_b200.sampling import Sample
    rng = np.random.default_rng(5)
    n_sites, n_features, n_categories = 10, 5, 3
    features = np.eye(n_categories, dtype=bool)[rng.choice(n_categories, size=(n_sites, n_features),
                                                           p=rng.dirichlet(np.ones(n_categories)))]
    areas = np.array([[1, 1, 1, 1, 0, 0, 0, 0, 0, 0]], dtype=bool)
    families = areas.copy()
    p_global = rng.dirichlet(np.ones(n_categories), size=(1, n_features))
    p_areas = rng.dirichlet(np.ones(n_categories), size=(1, n_features))
    p_families = p_areas.copy()
    w_with = np.repeat([[0.4, 0.3, 0.3]], n_features, axis=0)
    w_without = np.repeat([[0.4, 0.6]], n_features, axis=0)
    sample_with_family = Sample(areas, w_with, p_global=p_global, p_zones=p_areas, p_families=p_families)
    sample_without_family = Sample(areas, w_without, p_global=p_global, p_zones=p_areas, p_families=None)
    Data = namedtuple('Data', ['features', 'families'])
    data = Data(features=features, families=families)
    lh_with_family = Likelihood(data=data, inheritance=True)(sample_with_family, caching=False)
    lh_without_family = Likelihood(data=data, inheritance=False)(sample_without_family, caching=False)
    assert abs(lh_with_family - lh_without_family) < 1e-7           # assertAlmostEqual: 7 places
    p_mixed = 0.4 * p_global + 0.6 * p_areas
    p_per_feature = np.repeat(p_global, n_sites, axis=0)
    p_per_feature[areas[0]] = np.repeat(p_mixed, np.count_nonzero(areas), axis=0)
    lh_direct = np.sum(np.log(p_per_feature[features]))
    assert abs(lh_with_family - lh_direct) < 1e-7
    assert isinstance(lh_with_family, float)
    assert not sample_with_family.what_changed['lh']['zones']        # flags cleared like model.py:186-192


@pytest.mark.parametrize('name', golden_cases())
def test_model_objects_match_reference(name, built_lib):
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import Sample
    c = load_case(name)
    g = c['g']
    data = case_data(c)
    model = Model(data, case_model_config(c))
    with_src = Sample(g['zones'], g['weights'], g['p_global'], g['p_zones'], c['p_families'], source=c['source'])
    no_src = Sample(g['zones'], g['weights'], g['p_global'], g['p_zones'], c['p_families'])
    assert close(model.likelihood(with_src), float(g['ref_ll_src']))
    assert close(model.likelihood(no_src, caching=False), float(g['ref_ll_nosrc']))
    assert close(model.prior(no_src), float(g['ref_prior']))
    assert close(model(no_src), float(g['ref_ll_nosrc']) + float(g['ref_prior']))
    # the prior terms on their own add up to the joint prior; the geo term is the reference's (quirk Q1)
    pr = model.prior
    parts = pr.size_prior(no_src) + pr.geo_prior(no_src) + pr.prior_weights(no_src) + pr.prior_p_global(no_src) + \
        pr.prior_p_zones(no_src)
    if c['inheritance']:
        parts += pr.prior_p_families(no_src)
    assert close(parts, float(g['ref_prior']))
    assert close(pr.geo_prior(no_src), float(g['ref_geo']))
    # per-cell arrays of the operators' API
    assert np.max(np.abs(model.likelihood.update_weights(no_src) - g['ref_wnorm'])) < 1e-15
    assert model.likelihood.update_component_likelihoods(no_src).shape == (c['N'], c['F'], c['C'])
    assert np.array_equal(model.likelihood.get_zone_assignment(no_src), g['zones'].any(axis=0))


def test_prior_errors_like_reference(built_lib):
    from contact_zones_b200.model import GeoPrior, Model, PGlobalPrior, ZoneSizePrior
    from contact_zones_b200.sampling import Sample
    c = load_case(golden_cases()[0])
    data = case_data(c)
    with pytest.raises(ValueError):
        ZoneSizePrior({'type': 'cubic'}, data)
    with pytest.raises(ValueError):
        GeoPrior({'type': 'gaussian'}, data)
    with pytest.raises(ValueError):
        PGlobalPrior({'type': 'jeffrey'}, data)
    cfg = case_model_config(c)
    cfg['PRIOR']['contact'] = {'type': 'universal', 'scale_counts': 1.}
    g = c['g']
    with pytest.raises(ValueError):
        Model(data, cfg).prior(Sample(g['zones'], g['weights'], g['p_global'], g['p_zones'], c['p_families']))
