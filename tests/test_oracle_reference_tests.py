"""CPU: the reference's own known answers for this path, restated against the oracle.
  * test/test_model.py:52-92  (family == zone overlap identity + direct formula)
  * util.py:1212-1215 log_binom doctest, util.py:1098-1103 normalize doctest"""
import numpy as np

import sbayes_oracle as orc


def test_log_binom_doctest():
    assert np.allclose(orc.log_binom(10, np.arange(3)), [0., 2.30258509, 3.80666249], atol=1e-8)
    assert np.allclose(orc.log_binom(np.arange(1, 4), 1), [0., 0.69314718, 1.09861229], atol=1e-8)


def test_normalize_doctest():
    assert np.array_equal(orc.normalize(np.ones((2, 4))), np.full((2, 4), 0.25))
    assert np.array_equal(orc.normalize(np.ones((2, 4)), axis=0), np.full((2, 4), 0.5))


def test_family_area_overlap():
    """Restates test/test_model.py:52-92 with a seeded generator (the reference's test is unseeded)."""
    rng = np.random.default_rng(0)
    n, f, s = 10, 5, 3
    x = rng.integers(0, s, size=(n, f))
    features = np.eye(s, dtype=bool)[x]
    zones = np.zeros((1, n), bool)
    zones[0, :4] = True
    families = zones.copy()
    p_global = rng.dirichlet(np.ones(s), size=(1, f))
    p_zones = rng.dirichlet(np.ones(s), size=(1, f))
    w3 = np.tile([0.4, 0.3, 0.3], (f, 1))
    w2 = np.tile([0.4, 0.6], (f, 1))
    lh3 = orc.log_likelihood(features, families, True, zones, w3, p_global, p_zones, p_zones.copy())
    lh2 = orc.log_likelihood(features, families, False, zones, w2, p_global, p_zones, None)
    assert abs(lh3 - lh2) < 1e-7
    p_mix = np.tile(p_global, (n, 1, 1))
    p_mix[:4] = 0.4 * p_global + 0.6 * p_zones
    direct = np.sum(np.log(p_mix[features]))
    assert abs(lh2 - direct) < 1e-7
