"""GPU: the CUDA path (through the C ABI) against the golden vectors from the reference and against the
oracle on the same seeded inputs.

Bars (BASELINE.json north_star): log-likelihood / log-prior / geo-prior within 1e-9 relative (fp64), the
integer count tables bit-exact, per-cell arrays 1e-12 absolute."""
import numpy as np
import pytest

import sbayes_oracle as orc
from conftest import golden_cases
from helpers import load_case, make_engine, rel_err, set_golden_state

pytestmark = pytest.mark.gpu
RTOL = 1e-9


@pytest.mark.parametrize('name', golden_cases())
def test_eval_matches_reference(name, built_lib):
    c = load_case(name)
    g = c['g']
    eng = make_engine(c, n_chains=2, sample_source=True)
    for ch in range(2):
        set_golden_state(eng, c, ch)
    ll, prior, geo = eng.eval(use_source=True)
    assert rel_err(ll, g['ref_ll_src']) < RTOL
    assert abs(prior[0] - g['ref_prior']) <= RTOL * max(1.0, abs(g['ref_prior']))
    assert abs(geo[1] - g['ref_geo']) <= RTOL * max(1.0, abs(g['ref_geo']))
    ll0, _, _ = eng.eval(use_source=False)
    assert rel_err(ll0, g['ref_ll_nosrc']) < RTOL
    eng.close()
    # a context without source state evaluates the source-free likelihood
    eng = make_engine(c, n_chains=1, sample_source=False)
    set_golden_state(eng, c, 0, with_source=False)
    ll1, prior1, _ = eng.eval()
    assert rel_err(ll1, g['ref_ll_nosrc']) < RTOL
    assert abs(prior1[0] - g['ref_prior']) <= RTOL * max(1.0, abs(g['ref_prior']))
    eng.close()


@pytest.mark.parametrize('name', golden_cases())
def test_components_and_posterior(name, built_lib):
    c = load_case(name)
    g = c['g']
    eng = make_engine(c)
    set_golden_state(eng, c)
    lh, wn = eng.components(0)
    assert np.max(np.abs(wn - g['ref_wnorm'])) < 1e-15
    ref_lh = orc.component_likelihoods(c['features'], g['zones'], c['families'], g['p_global'], g['p_zones'],
                                       c['p_families'], c['inheritance'])
    assert np.max(np.abs(lh - ref_lh)) == 0.0      # pure table look-ups
    assert np.max(np.abs(eng.source_posterior(0) - g['ref_post'])) < 1e-14
    eng.close()


@pytest.mark.parametrize('name', golden_cases())
def test_counts_bit_exact(name, built_lib):
    c = load_case(name)
    g = c['g']
    eng = make_engine(c)
    set_golden_state(eng, c)
    for recount in (False, True):
        cnt = eng.counts(0, recount=recount)
        assert np.array_equal(cnt['cnt_glob'], g['ref_cnt_glob'])
        assert np.array_equal(cnt['cnt_zone'], g['ref_cnt_zone'])
        assert np.array_equal(cnt['cnt_src_zone'], g['ref_cnt_src_zone'])
        if c['inheritance']:
            assert np.array_equal(cnt['cnt_fam'], g['ref_cnt_fam'])
            assert np.array_equal(cnt['cnt_src_fam'], g['ref_cnt_src_fam'])
    eng.close()


@pytest.mark.parametrize('name', golden_cases())
def test_move_probabilities(name, built_lib):
    c = load_case(name)
    g = c['g']
    eng = make_engine(c)
    set_golden_state(eng, c)
    for kind, z, sn, sr, lq, lqb in g['ref_moves']:
        k = ['grow', 'shrink', 'swap'][int(kind)]
        q, qb = eng.move_log_q(0, k, int(z), int(sn), int(sr), 0.85)
        assert abs(q - lq) < 1e-12 and abs(qb - lqb) < 1e-12, (k, q, lq, qb, lqb)
    eng.close()


def test_state_roundtrip_and_overlap(built_lib):
    from contact_zones_b200._lib import SbError
    c = load_case('D_small')
    g = c['g']
    eng = make_engine(c, n_chains=2)
    set_golden_state(eng, c, 1)
    s = eng.get_state(1)
    assert np.array_equal(s['zones'], g['zones'])
    assert np.array_equal(s['weights'], g['weights']) and np.array_equal(s['p_zones'], g['p_zones'])
    assert np.array_equal(s['p_global'][0], g['p_global'][0]) and np.array_equal(s['p_families'], g['p_families'])
    assert np.array_equal(np.argmax(s['source'], -1), g['source_idx'])
    eng.copy_state(0, 1)
    assert np.array_equal(eng.get_state(0)['zones'], g['zones'])
    bad = g['zones'].copy()
    bad[1] |= bad[0]
    with pytest.raises(SbError):
        eng.set_state(0, bad, g['weights'], g['p_global'], g['p_zones'], c['p_families'], g['source_idx'])
    eng.close()


def test_edge_cases(built_lib):
    """-inf for a source pointing at a zero-weight component (model.py:181-182); zero probability -> -inf;
    NA-only rows; a zone-free state; tiny probabilities (slow path of the product accumulator)."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine
    data = synthetic.make_data(64, 9, 3, n_families=0, na_frac=0.3, seed=5)
    data.features[3] = False                      # a site with only NA cells
    st = synthetic.random_state(data, 2, 4, False, True, seed=6)
    eng = Engine(data.features, data.states, None, data.network['adj_mat'], None, n_zones=2, n_chains=1,
                 inheritance=False, sample_source=True)
    args = (data.features, data.families, False, st['zones'], st['weights'], st['p_global'], st['p_zones'], None)

    def both(src, **over):
        a = list(args)
        for k, v in over.items():
            a[{'zones': 3, 'weights': 4, 'p_global': 5, 'p_zones': 6}[k]] = v
        eng.set_state(0, a[3], a[4], a[5], a[6], None, src)
        return (eng.eval(use_source=True)[0][0], eng.eval(use_source=False)[0][0],
                orc.log_likelihood(*a, source=src), orc.log_likelihood(*a, source=None))

    g1, g0, o1, o0 = both(st['source'])
    assert rel_err(g1, o1) < RTOL and rel_err(g0, o0) < RTOL
    # source -> contact at a site that is in no zone: chosen weight is 0 -> -inf
    src = st['source'].copy()
    free = np.flatnonzero(~st['zones'].any(0))[0]
    src[free, 0] = [False, True]
    g1, g0, o1, o0 = both(src)
    assert g1 == -np.inf and o1 == -np.inf and rel_err(g0, o0) < RTOL
    # zero and tiny probabilities
    pg = st['p_global'].copy()
    pg[0, 0, :] = [1e-250, 1.0 - 1e-250, 0.0][:pg.shape[2]] if pg.shape[2] == 3 else pg[0, 0, :]
    g1, g0, o1, o0 = both(st['source'], p_global=pg)
    assert (g0 == o0) or rel_err(g0, o0) < RTOL
    assert (g1 == o1) or rel_err(g1, o1) < RTOL
    # no site in any zone
    z0 = np.zeros_like(st['zones'])
    src0 = np.zeros_like(st['source'])
    src0[..., 0] = True
    g1, g0, o1, o0 = both(src0, zones=z0)
    assert rel_err(g1, o1) < RTOL and rel_err(g0, o0) < RTOL
    eng.close()


def test_large_shape_properties(built_lib):
    """BASELINE.json shape D (2000 x 200 x 8, 7 zones, 20 families): oracle parity plus size-independent
    properties: chains holding the same state give the same bits; family == zone overlap identity
    (test/test_model.py:52-92) at full size."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine
    data = synthetic.make_data(2000, 200, 8, n_families=20, seed=0, with_dist=False)
    st = synthetic.random_state(data, 7, 20, True, True, seed=1)
    eng = Engine(data.features, data.states, data.families, data.network['adj_mat'], None, n_zones=7, n_chains=3,
                 inheritance=True, sample_source=True)
    for ch in range(3):
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'], st['source'])
    fam = np.asarray(data.families, bool)
    args = (data.features, fam, True, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'])
    ll1 = eng.eval(use_source=True)[0]
    ll0 = eng.eval(use_source=False)[0]
    assert ll1[0] == ll1[1] == ll1[2] and ll0[0] == ll0[1] == ll0[2]
    assert rel_err(ll1[0], orc.log_likelihood(*args, source=st['source'])) < RTOL
    assert rel_err(ll0[0], orc.log_likelihood(*args, source=None)) < RTOL
    cnt = eng.counts(0, recount=True)
    assert np.array_equal(cnt['cnt_glob'], orc.state_counts_global(data.features, st['source']))
    assert np.array_equal(cnt['cnt_fam'][7], orc.state_counts_family(data.features, st['source'], fam[7]))
    assert np.array_equal(cnt['cnt_src_fam'], orc.source_counts(st['source'], fam.any(0)))
    eng.close()


def test_bulk_upload_equals_single_uploads(built_lib):
    """sb_set_states (arrays stacked along the chain axis) == sb_set_state chain by chain: identical evaluation,
    count tables and read-back."""
    from contact_zones_b200 import synthetic
    c = load_case('C_samerica_like')
    g = c['g']
    data_ns = type('D', (), {})()
    rng = np.random.default_rng(4)
    n = 5
    zones, src = [], []
    for i in range(n):
        z = np.zeros_like(g['zones'])
        perm = rng.permutation(c['N'])
        for k in range(c['Z']):
            z[k, perm[k * 6:(k + 1) * 6]] = True
        zones.append(z)
        s = rng.integers(0, c['C'], size=(c['N'], c['F'])).astype(np.uint8)
        s[~z.any(axis=0)] = np.where(s[~z.any(axis=0)] == 1, 0, s[~z.any(axis=0)])          # no zone -> not contact
        s[~c['families'].any(axis=0)] = np.where(s[~c['families'].any(axis=0)] == 2, 0, s[~c['families'].any(axis=0)])
        src.append(s)
    w = rng.dirichlet(np.ones(c['C']), size=(n, c['F']))
    pg = np.stack([g['p_global'][0]] * n)
    pz = np.stack([g['p_zones'][rng.permutation(c['Z'])] for _ in range(n)])
    pf = np.stack([c['p_families']] * n)
    a = make_engine(c, n_chains=n)
    b = make_engine(c, n_chains=n)
    a.set_states(0, np.stack(zones), w, pg, pz, pf, np.stack(src))
    for i in range(n):
        b.set_state(i, zones[i], w[i], pg[i], pz[i], pf[i], src[i])
    ra, rb = a.eval(), b.eval()
    for x, y in zip(ra, rb):
        assert np.array_equal(x, y)
    assert len(set(np.round(ra[0], 6))) == n                       # the chains really differ
    for i in (0, n - 1):
        ca, cb = a.counts(i), b.counts(i)
        for k in ca:
            assert np.array_equal(ca[k], cb[k]), k
        sa, sb = a.get_state(i), b.get_state(i)
        for k in sa:
            assert np.array_equal(sa[k], sb[k]), k
        assert np.array_equal(sa['zones'], zones[i]) and np.array_equal(np.argmax(sa['source'], -1), src[i])
    # page-locked caller memory takes the unstaged path (direct copies, pitched for the source rows: F = 36 is not a
    # multiple of 16); an out-of-range source value must still be refused and leave the call clean
    import torch
    pin = [torch.from_numpy(np.ascontiguousarray(v)).pin_memory().numpy()
           for v in (np.stack(zones).astype(np.uint8), w, pg, pz, pf, np.stack(src))]
    p = make_engine(c, n_chains=n)
    p.set_states(0, *pin)
    for x, y in zip(p.eval(), ra):
        assert np.array_equal(x, y)
    for i in (0, n - 1):
        cp, ca = p.counts(i), a.counts(i)
        for k in cp:
            assert np.array_equal(cp[k], ca[k]), k
        sp_, sa = p.get_state(i), a.get_state(i)
        for k in sp_:
            assert np.array_equal(sp_[k], sa[k]), k
    pin[5][2, 3, 4] = 7
    with pytest.raises(Exception):
        p.set_states(0, *pin)
    pin[5][2, 3, 4] = src[2][3, 4]
    p.set_states(0, *pin)
    for x, y in zip(p.eval(), ra):
        assert np.array_equal(x, y)
    a.close(); b.close(); p.close()


@pytest.mark.parametrize('name', ['C_samerica_like', 'G_samerica_data'])
def test_fallback_evaluation_kernels(name, built_lib, monkeypatch):
    """SBAYES_B200_SWEEP=0 selects the tile-per-site fallbacks (count_kernel, lh_full_kernel<1>: used when the tables of
    a 16-feature block do not fit in shared memory): same likelihood within 1e-9, same count tables bit for bit."""
    c = load_case(name)
    g = c['g']
    monkeypatch.setenv('SBAYES_B200_SWEEP', '0')
    eng = make_engine(c)
    set_golden_state(eng, c)
    ll, _, _ = eng.eval()
    assert rel_err(ll[0], float(g['ref_ll_src'])) < 1e-9
    slow = eng.counts(0, recount=True)
    eng.close()
    monkeypatch.delenv('SBAYES_B200_SWEEP')
    eng = make_engine(c)
    set_golden_state(eng, c)
    ll2, _, _ = eng.eval()
    fast = eng.counts(0, recount=True)
    eng.close()
    assert rel_err(ll2[0], ll[0]) < 1e-12
    for k in fast:
        if fast[k] is not None:
            assert np.array_equal(fast[k], slow[k]), k


@pytest.mark.parametrize('max_n', ['0', '1000000', 'chain'])
@pytest.mark.parametrize('name', golden_cases())
def test_both_source_free_kernels_on_the_goldens(name, max_n, built_lib, monkeypatch):
    """The source-free likelihood has three kernels: the table kernel (packed state codes, sb_lhtab.cuh), its
    one-CTA-per-chain form for many chains over few sites, and the blocked kernel (one CTA per range of 16-feature blocks,
    sb_eval.cuh), chosen by shape.  SB_K1_BLOCK_MAX_N = 0 keeps every shape with two or more sites per code byte on the
    table kernel, 1000000 sends every shape to the blocked one (with 1 .. all blocks per CTA, SB_K1B_PARTS), 'chain'
    sends the shapes with two or more sites per code byte to the one-CTA-per-chain kernel: all reproduce the
    reference's value."""
    c = load_case(name)
    g = c['g']
    monkeypatch.setenv('SB_K1C_MIN_CHAINS', '1' if max_n == 'chain' else '1000000')
    if max_n == 'chain':
        max_n = '0'
    monkeypatch.setenv('SB_K1_BLOCK_MAX_N', max_n)
    for parts in (['0'] if max_n == '0' else ['0', '1', '2', '1000']):
        monkeypatch.setenv('SB_K1B_PARTS', parts)
        eng = make_engine(c, n_chains=3, sample_source=False)
        for ch in range(3):
            set_golden_state(eng, c, ch, with_source=False)
        ll = eng.eval()[0]
        assert rel_err(ll, g['ref_ll_nosrc']) < RTOL, (parts, ll)
        assert ll[0] == ll[1] == ll[2]
        eng.close()


@pytest.mark.parametrize('seed', range(int(__import__('os').environ.get('SB_FUZZ_SEEDS_K1', '10'))))
def test_likelihood_kernels_on_random_large_shapes(seed, built_lib):
    """K1 (table-driven, source-free), K1 with source and the K4 recount on seeded random shapes up to 20000 sites,
    40 families, 8 zones, 1..300 chains (the chain count drives K1's split into feature chunks / group splits):
    likelihood == oracle within 1e-9, count tables == oracle bit for bit."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    rng = np.random.default_rng(7000 + seed)
    n = int(rng.choice([40, 333, 1000, 4097, 9000, 20000]))
    f, s_ = int(rng.integers(1, 50)), int(rng.integers(2, 9))
    fam = int(rng.integers(0, 41))
    z, zs = int(rng.integers(1, 9)), int(rng.integers(3, 30))
    inh = fam > 0
    n_chains = int(rng.choice([1, 2, 7, 64, 300]))
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=seed, with_dist=False, ragged_states=bool(rng.integers(0, 2)),
                               na_frac=float(rng.choice([0.0, 0.02, 0.5])))
    try:
        sts = [synthetic.random_state(data, z, min(zs, max(3, n // (2 * z))), inh, True, seed=seed * 7 + i) for i in range(2)]
    except ValueError:
        pytest.skip('could not place the random zones')
    eng = Engine(state_index(data.features), data.states, data.families if inh else None, data.network['adj_mat'], None,
                 n_zones=z, n_chains=n_chains, inheritance=inh, sample_source=True)
    for ch in range(n_chains):
        st = sts[ch % 2]
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'] if inh else None,
                      np.argmax(st['source'], -1))
    ll_free = eng.eval(use_source=False)[0]
    ll_src = eng.eval(use_source=True)[0]
    fam_b = np.asarray(data.families, bool)
    for ch in sorted({0, n_chains - 1, min(1, n_chains - 1)}):
        st = sts[ch % 2]
        args = (data.features, fam_b, inh, st['zones'], st['weights'], st['p_global'], st['p_zones'],
                st['p_families'] if inh else None)
        ref = orc.log_likelihood(*args, source=None)
        assert abs(ll_free[ch] - ref) <= RTOL * abs(ref), (ch, ll_free[ch], ref)
        ref = orc.log_likelihood(*args, source=st['source'])
        assert abs(ll_src[ch] - ref) <= RTOL * abs(ref), (ch, ll_src[ch], ref)
    st = sts[(n_chains - 1) % 2]
    cnt = eng.counts(n_chains - 1, recount=True)
    assert np.array_equal(cnt['cnt_glob'], orc.state_counts_global(data.features, st['source']))
    for k in range(z):
        assert np.array_equal(cnt['cnt_zone'][k], orc.state_counts_zone(data.features, st['source'], st['zones'][k]))
    eng.close()


@pytest.mark.parametrize('shape', [(40000, 5, 3, 2, 3), (300, 20, 32, 3, 5), (500, 17, 8, 2, 150), (1500, 33, 8, 4, 0)])
def test_blocked_kernels_at_the_edges_of_their_shapes(shape, built_lib):
    """The blocked evaluation kernels (K1b, K1s, K4) where their layout changes: more than 32768 sites per chain (K4's
    per-thread site tallies are flushed early), S = 32 (33 state columns, no table kernel), 150 families (the tables
    of a block exceed shared memory: tile-per-site fallbacks), no families at S = 8."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    n, f, s_, z, fam = shape
    inh = fam > 0
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=n + f, with_dist=False, na_frac=0.05)
    sts = [synthetic.random_state(data, z, 12, inh, True, seed=50 + i) for i in range(2)]
    eng = Engine(state_index(data.features), data.states, data.families if inh else None, data.network['adj_mat'], None,
                 n_zones=z, n_chains=3, inheritance=inh, sample_source=True)
    for ch in range(3):
        st = sts[ch % 2]
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'] if inh else None,
                      np.argmax(st['source'], -1))
    ll_free, ll_src = eng.eval(use_source=False)[0], eng.eval(use_source=True)[0]
    fam_b = np.asarray(data.families, bool)
    for ch in (0, 1):
        st = sts[ch]
        args = (data.features, fam_b, inh, st['zones'], st['weights'], st['p_global'], st['p_zones'],
                st['p_families'] if inh else None)
        assert rel_err(ll_free[ch], orc.log_likelihood(*args, source=None)) < RTOL
        assert rel_err(ll_src[ch], orc.log_likelihood(*args, source=st['source'])) < RTOL
    assert ll_free[2] == ll_free[0] and ll_src[2] == ll_src[0]
    st = sts[1]
    cnt = eng.counts(1, recount=True)
    assert np.array_equal(cnt['cnt_glob'], orc.state_counts_global(data.features, st['source']))
    for k in range(z):
        assert np.array_equal(cnt['cnt_zone'][k], orc.state_counts_zone(data.features, st['source'], st['zones'][k]))
    # the source tables of gibbs_sample_weights (derived inside K4 from the per-configuration global tables)
    assert np.array_equal(cnt['cnt_src_zone'], orc.source_counts(st['source'], st['zones'].any(0)))
    if inh:
        assert np.array_equal(cnt['cnt_src_fam'], orc.source_counts(st['source'], fam_b.any(0)))
    eng.close()
