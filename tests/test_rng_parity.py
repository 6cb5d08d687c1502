"""GPU: the device random-number layer K6 (Philox4x32-10 + Marsaglia-Tsang Gamma) against the distributions the
reference draws from numpy / scipy (np.random.dirichlet -> normalised Gammas, zone_sampling.py:243,351,373,401,560;
scipy.stats.beta.rvs :262-265 -> Gamma ratio).  Kolmogorov-Smirnov against the exact CDFs plus moment checks;
p-value bar 1e-4 per test (fixed seeds, deterministic)."""
import numpy as np
import pytest
from scipy import stats

from helpers import load_case, make_engine

pytestmark = pytest.mark.gpu
N = 200_000


@pytest.fixture(scope='module')
def eng(built_lib):
    e = make_engine(load_case('E_small'), sample_source=False)
    yield e
    e.close()


def test_uniform(eng):
    u = eng.rng_sample('uniform', N, seed=1)
    assert 0.0 < u.min() and u.max() < 1.0
    assert stats.kstest(u, 'uniform').pvalue > 1e-4
    assert abs(np.corrcoef(u[:-1], u[1:])[0, 1]) < 0.01
    assert not np.array_equal(u, eng.rng_sample('uniform', N, seed=2))
    assert np.array_equal(u, eng.rng_sample('uniform', N, seed=1))           # counter-based: reproducible


@pytest.mark.parametrize('alpha', [0.3, 1.0, 1.7, 16.0, 401.0, 25000.0])
def test_gamma(eng, alpha):
    g = eng.rng_sample('gamma', N, alpha=alpha, seed=int(alpha * 10) + 3)
    assert g.min() > 0.0
    assert stats.kstest(g, 'gamma', args=(alpha,)).pvalue > 1e-4
    se = np.sqrt(alpha / N)
    assert abs(g.mean() - alpha) < 5 * se
    assert abs(g.var() - alpha) < 5 * alpha * np.sqrt((2 + 6 / alpha) / N)


def test_normal_candidates(eng):
    """The N(0,1) candidates of the Gamma sampler come from hardware-approximate lg2 / sqrt / sin / cos: the
    perturbation (~5e-7 absolute) must be invisible at 2e6 draws, in the body, the tails and the pair structure."""
    n = 2_000_000
    x = eng.rng_sample('normal', n, seed=5)
    assert stats.kstest(x, 'norm').pvalue > 1e-4
    assert abs(x.mean()) < 5 / np.sqrt(n) and abs(x.var() - 1) < 5 * np.sqrt(2 / n)
    assert abs(np.mean(x ** 4) - 3) < 5 * np.sqrt(96 / n)
    for t in (2.0, 3.0, 4.0):
        p = 2 * stats.norm.sf(t)
        assert abs(np.mean(np.abs(x) > t) - p) < 5 * np.sqrt(p / n)
    assert abs(np.corrcoef(x[0::2], x[1::2])[0, 1]) < 5 / np.sqrt(n / 2)      # the two normals of a pair
    r2 = x[0::2] ** 2 + x[1::2] ** 2                                           # radius^2 ~ chi2(2) = Exp(1/2)
    assert stats.kstest(r2, 'expon', args=(0, 2)).pvalue > 1e-4
    ang = np.arctan2(x[1::2], x[0::2])
    assert stats.kstest(ang, 'uniform', args=(-np.pi, 2 * np.pi)).pvalue > 1e-4


def test_beta_and_dirichlet_from_gammas(eng):
    """Beta(a, b) = Ga / (Ga + Gb) (gibbs_sample_weights) and a Dirichlet row (gibbs_sample_p_*)."""
    a, b = 3.0, 11.0
    ga, gb = eng.rng_sample('gamma', N, alpha=a, seed=11), eng.rng_sample('gamma', N, alpha=b, seed=12)
    assert stats.kstest(ga / (ga + gb), 'beta', args=(a, b)).pvalue > 1e-4
    alphas = np.array([1.0, 4.0, 2.5, 9.0])
    g = np.stack([eng.rng_sample('gamma', N, alpha=v, seed=20 + i) for i, v in enumerate(alphas)], axis=1)
    p = g / g.sum(axis=1, keepdims=True)
    assert np.allclose(p.mean(axis=0), alphas / alphas.sum(), atol=4e-3)
    a0 = alphas.sum()
    assert np.allclose(p.var(axis=0), alphas * (a0 - alphas) / (a0 ** 2 * (a0 + 1)), rtol=0.05)


def test_table_log_accuracy(eng):
    """flog() of the step kernel (sb_common.cuh) against numpy over 120 binades: absolute error <= 4e-16 + one ulp of the
    result (the final e*ln2 + ... fma rounds once at the magnitude of the result)."""
    n = 121 * 2000
    u = eng.rng_sample('uniform', n, seed=9)
    x = np.ldexp(1.0 + u, np.arange(n) % 121 - 60)
    got = eng.rng_sample('flog', n, seed=9)
    ref = np.log(x)
    err = np.abs(got - ref)
    assert np.all(err <= 4e-16 + np.spacing(np.abs(ref))), float(np.max(err - np.spacing(np.abs(ref))))
