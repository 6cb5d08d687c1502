"""CPU: host logic around the hot path -- label matching against the reference's brute-force definition
(postprocessing.py:205-236), ranking, the area bit-string codec (util.py:64-86), Sample / IndexSet semantics
(zone_sampling.py:16-115), the warm-up size ladder (util.py:1182-1199) and the operator table (mcmc_setup.py:70-95)."""
from itertools import permutations

import numpy as np

from contact_zones_b200 import results
from contact_zones_b200.sampling import IndexSet, Sample, get_max_size_list


def _brute_force_matching(sample_zones):
    """Literal restatement of postprocessing.py:214-236."""
    a = np.swapaxes(np.array([np.array(s) for s in sample_zones]), 1, 2)
    _, n_sites, n_areas = a.shape
    s_sum = np.zeros((n_sites, n_areas))
    perm = list(permutations(range(n_areas)))
    out = []
    for s in a:
        best = max(perm, key=lambda p: np.sum(s_sum * s[:, p]))
        out.append(list(best))
        s_sum += s[:, best]
    return out


def test_match_areas_equals_brute_force():
    rng = np.random.default_rng(3)
    n_samples, z, n = 25, 4, 60
    base = np.zeros((z, n), bool)
    for k in range(z):
        base[k, k * 12:(k + 1) * 12] = True
    zones, pz, lh = [], [], []
    for _ in range(n_samples):
        p = rng.permutation(z)
        s = base[p].copy()
        flip = rng.integers(0, n, 4)
        s[:, flip] = False
        zones.append(s)
        pz.append(rng.random((z, 3, 2)))
        lh.append(list(rng.random(z)))
    want = _brute_force_matching(zones)
    samples = {'sample_zones': [x.copy() for x in zones], 'sample_p_zones': [x.copy() for x in pz],
               'sample_lh_single_zones': [list(x) for x in lh], 'sample_prior_single_zones': [list(x) for x in lh],
               'sample_posterior_single_zones': [list(x) for x in lh]}
    got = results.match_areas(samples)
    for s in range(n_samples):
        assert np.array_equal(got['sample_zones'][s], zones[s][want[s]])
        assert np.array_equal(got['sample_p_zones'][s], pz[s][want[s]])
        assert got['sample_lh_single_zones'][s] == [lh[s][i] for i in want[s]]
    # after matching, the planted blocks line up across samples
    agree = np.mean([np.array_equal(np.argmax(g @ base.T, axis=1), np.argmax(got['sample_zones'][0] @ base.T, axis=1))
                     for g in got['sample_zones']])
    assert agree == 1.0
    ranked = results.rank_areas(got)
    mean_post = np.mean(np.asarray(ranked['sample_posterior_single_zones']), axis=0)
    assert np.all(np.diff(mean_post) <= 0)


def test_area_codec_round_trip():
    rng = np.random.default_rng(0)
    z = rng.random((3, 17)) < 0.3
    line = results.format_area_columns(z)
    assert line.count('\t') == 2 and set(line) <= {'0', '1', '\t'}
    assert np.array_equal(results.parse_area_columns(line), z)


def test_sample_and_indexset_semantics():
    s = Sample(np.zeros((1, 4), bool), np.ones((2, 2)) / 2, np.ones((1, 2, 2)) / 2, np.ones((1, 2, 2)) / 2, None)
    assert s.what_changed['lh']['zones'] and s.what_changed['prior']['weights'] is True
    s.what_changed['lh']['zones'].clear()
    assert not s.what_changed['lh']['zones']
    s.what_changed['lh']['zones'].add(0)
    t = s.copy()
    t.zones[0, 0] = True
    t.what_changed['lh']['zones'].add(1)
    assert not s.zones[0, 0] and s.what_changed['lh']['zones'] == {0} and t.what_changed['lh']['zones'] == {0, 1}
    e = Sample.empty_sample()
    assert e.zones is None and e.what_changed['lh']['p_global']
    i = IndexSet(all_i=False)
    assert not i
    i.add(3)
    assert i and 3 in i


def test_max_size_list_matches_reference_formula():
    got = get_max_size_list(start=(5 + 50) / 4, end=50, n_total=15, k_groups=4)
    assert got == [13] * 4 + [22] * 4 + [31] * 4 + [40] * 3
    assert len(get_max_size_list(10, 30, 3, 4)) == 3


def test_sub_seeds_differ_per_sampler_instance():
    """The config's SEED is expanded into one sub-seed per (run, N_AREAS, phase): warm-up and sampling, the runs of
    N_RUNS and the values of an N_AREAS list never replay each other's device streams."""
    from contact_zones_b200.mcmc_setup import MCMC
    seen = {MCMC.sub_seed(7, run, z, phase) for run in range(4) for z in range(1, 5) for phase in (0, 1)}
    assert len(seen) == 32 and all(0 <= v < 2 ** 63 for v in seen)
    assert MCMC.sub_seed(7, 1, 3, 1) == MCMC.sub_seed(7, 1, 3, 1) != MCMC.sub_seed(8, 1, 3, 1)


def test_overlapping_families_last_one_wins():
    """A site listed in two (nested) families belongs to the later one, like the reference's
    compute_family_likelihood (model.py:427-431), instead of being rejected."""
    from contact_zones_b200.engine import family_index
    fam = np.zeros((3, 6), bool)
    fam[0, :4] = True
    fam[1, 2:5] = True
    assert list(family_index(fam, 6)) == [0, 0, 1, 1, 1, -1]
