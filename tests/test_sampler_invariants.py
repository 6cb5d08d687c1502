"""GPU: structural invariants of the step kernel K8 through the C ABI.

* every operator on its own keeps the incrementally maintained log-likelihood / log-prior / count tables equal to a
  from-scratch evaluation (K1, K2, K3, K4) and the zone book-keeping consistent (disjoint zones, MIN_M <= size <= MAX_M);
* a chain's trajectory depends on (seed, global chain id) only -- not on how chains are sharded over contexts;
* teams of 1, 2, 4 and 8 CTAs per chain (thread-block clusters) walk the same trajectory: the team size only changes the
  order of fp64 partial sums (1e-16 relative), which cannot flip an accept decision within a few hundred steps."""
import os

import numpy as np
import pytest

from helpers import PRECISION, load_case, make_engine, set_golden_state

pytestmark = pytest.mark.gpu

GIBBS = ['gibbs_sample_sources', 'gibbs_sample_weights', 'gibbs_sample_p_global', 'gibbs_sample_p_zones',
         'gibbs_sample_p_families']
ALTER = ['alter_weights', 'alter_p_global', 'alter_p_zones', 'alter_p_families']
ZONE = ['grow_zone', 'shrink_zone', 'swap_zone']


def check_consistency(eng, c, n_chains, sample_source, min_size, max_size):
    ll_run, pr_run = eng.chain_scalars()
    cnt_run = [eng.counts(ch, recount=False) for ch in range(n_chains)] if sample_source else None
    ll_new, pr_new, _ = eng.eval()
    assert np.allclose(ll_run, ll_new, rtol=1e-9), (ll_run, ll_new)
    assert np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9), (pr_run, pr_new)
    for ch in range(n_chains):
        z = eng.get_state(ch, fields=('zones',))['zones']
        assert (z.sum(axis=0) <= 1).all()
        assert (z.sum(axis=1) >= min(min_size, c['g']['zones'].sum(1).min())).all() and (z.sum(axis=1) <= max_size).all()
        if sample_source:
            new = eng.counts(ch, recount=True)
            for k in new:
                if new[k] is not None:
                    assert np.array_equal(cnt_run[ch][k], new[k]), (ch, k)


@pytest.mark.parametrize('name', ['B_balkan_like', 'D_small'])
@pytest.mark.parametrize('sample_source', [True, False])
def test_each_operator_keeps_running_values_exact(name, sample_source, built_lib):
    c = load_case(name)
    n_chains = 3
    eng = make_engine(c, n_chains=n_chains, sample_source=sample_source)
    for ch in range(n_chains):
        set_golden_state(eng, c, ch, with_source=sample_source)
    eng.eval()
    ops = ZONE + (GIBBS if sample_source else ALTER)
    if not c['inheritance']:
        ops = [o for o in ops if 'families' not in o]
    for i, op in enumerate(ops):
        eng.set_operators({op: 1.0}, PRECISION, min_size=3, max_size=25, p_grow_connected=0.85, seed=100 + i)
        eng.run(40 if op == 'gibbs_sample_sources' else 300)
        check_consistency(eng, c, n_chains, sample_source, 3, 25)
    acc, rej = eng.op_stats()
    from contact_zones_b200 import _lib
    tot = acc + rej
    for op in ops:
        assert (tot[:, _lib.OP_IDS[op]] == (40 if op == 'gibbs_sample_sources' else 300)).all(), op
    for op in GIBBS:
        if op in ops:
            assert (rej[:, _lib.OP_IDS[op]] == 0).all()              # Gibbs operators are always accepted
    for op in ('grow_zone', 'swap_zone', 'alter_p_zones'):
        if op in ops:
            assert acc[:, _lib.OP_IDS[op]].sum() > 0, op                 # ... and the MH ones do move
    eng.close()


def _run_chains(c, n_chains, chain_offset, steps, ops, sample_source=True, team=None, sweep=None):
    old = os.environ.get('SBAYES_B200_TEAM')
    old_sweep = os.environ.get('SBAYES_B200_SWEEP')
    if team is not None:
        os.environ['SBAYES_B200_TEAM'] = str(team)
    if sweep is not None:
        os.environ['SBAYES_B200_SWEEP'] = str(sweep)
    try:
        eng = make_engine(c, n_chains=n_chains, sample_source=sample_source, chain_offset=chain_offset)
        for ch in range(n_chains):
            set_golden_state(eng, c, ch, with_source=sample_source)
        eng.eval()
        eng.set_operators(ops, PRECISION, min_size=3, max_size=25, p_grow_connected=0.85, seed=77)
        eng.run(steps)
        out = [eng.get_state(ch, one_hot_source=False) for ch in range(n_chains)]
        ll, pr = eng.chain_scalars()
        eng.close()
        return out, ll, pr
    finally:
        for key, was, given in (('SBAYES_B200_TEAM', old, team), ('SBAYES_B200_SWEEP', old_sweep, sweep)):
            if given is not None:
                if was is None:
                    os.environ.pop(key, None)
                else:
                    os.environ[key] = was


MIX = {'shrink_zone': 0.1, 'grow_zone': 0.1, 'swap_zone': 0.05, 'gibbs_sample_weights': 0.3, 'gibbs_sample_p_global': 0.1,
       'gibbs_sample_p_zones': 0.25, 'gibbs_sample_p_families': 0.1}


def test_trajectories_do_not_depend_on_sharding(built_lib):
    c = load_case('D_small')
    full, ll_f, _ = _run_chains(c, 4, 0, 400, MIX, team=2)
    part, ll_p, _ = _run_chains(c, 2, 2, 400, MIX, team=2)
    assert np.array_equal(ll_f[2:], ll_p)
    for a, b in zip(full[2:], part):
        for k in a:
            if a[k] is not None:
                assert np.array_equal(a[k], b[k]), k
    assert not np.array_equal(full[0]['zones'], full[1]['zones'])         # different chains do differ


@pytest.mark.parametrize('sample_source', [True, False])
def test_team_sizes_walk_the_same_trajectory(sample_source, built_lib):
    c = load_case('D_small')
    ops = MIX if sample_source else {'shrink_zone': 0.1, 'grow_zone': 0.1, 'swap_zone': 0.05, 'alter_weights': 0.3,
                                     'alter_p_global': 0.1, 'alter_p_zones': 0.25, 'alter_p_families': 0.1}
    ref, ll_ref, pr_ref = _run_chains(c, 2, 0, 300, ops, sample_source, team=1)
    for team in (2, 4, 8):
        got, ll, pr = _run_chains(c, 2, 0, 300, ops, sample_source, team=team)
        assert np.allclose(ll, ll_ref, rtol=1e-10) and np.allclose(pr, pr_ref, rtol=1e-10, atol=1e-10), team
        for a, b in zip(ref, got):
            assert np.array_equal(a['zones'], b['zones']), team
            if sample_source:
                assert np.array_equal(a['source'], b['source']), team
            assert np.allclose(a['p_zones'], b['p_zones'], rtol=1e-12, atol=0), team
            assert np.allclose(a['weights'], b['weights'], rtol=1e-12, atol=0), team


def test_source_sweep_variants_produce_the_same_states(built_lib):
    """The two shared-memory forms of the source sweep (counters only / counters + staged parameter tables) use the
    same Philox blocks and the same arithmetic per cell: identical zones, sources and count tables, parameters equal
    to rounding, for teams of 1 and 2 (SBAYES_B200_SWEEP caps the variant).  The global-atomics fallback (variant 0)
    draws its uniforms differently; it must keep the running values consistent with a from-scratch evaluation."""
    c = load_case('D_small')
    ops = dict(MIX, gibbs_sample_sources=0.1)
    ref, ll_ref, pr_ref = _run_chains(c, 2, 0, 300, ops, True, team=1, sweep=1)
    for team, sweep in ((2, 1), (1, 2), (2, 2)):
        got, ll, pr = _run_chains(c, 2, 0, 300, ops, True, team=team, sweep=sweep)
        assert np.allclose(ll, ll_ref, rtol=1e-10) and np.allclose(pr, pr_ref, rtol=1e-10, atol=1e-10), (team, sweep)
        for a, b in zip(ref, got):
            assert np.array_equal(a['zones'], b['zones']) and np.array_equal(a['source'], b['source']), (team, sweep)
            assert np.allclose(a['p_zones'], b['p_zones'], rtol=1e-12, atol=0), (team, sweep)
    os.environ['SBAYES_B200_SWEEP'] = '0'
    try:
        eng = make_engine(c, n_chains=2, sample_source=True)
        for ch in range(2):
            set_golden_state(eng, c, ch)
        eng.eval()
        eng.set_operators(ops, PRECISION, min_size=3, max_size=25, p_grow_connected=0.85, seed=77)
        eng.run(300)
        check_consistency(eng, c, 2, True, 3, 25)
        eng.close()
    finally:
        os.environ.pop('SBAYES_B200_SWEEP', None)


@pytest.mark.parametrize('sample_source', [True, False])
def test_zone_cache_does_not_change_the_trajectory(sample_source, built_lib, monkeypatch):
    """The proposal phase of the zone moves reads the chain's zone ids / member lists / sizes and the adjacency CSR from
    shared memory when they fit (every CTA of a team keeps its own copy); SBAYES_B200_ZONE_CACHE=0 reads the global
    arrays: identical states either way, for teams of 1 and 2, in a zone-move-heavy mix."""
    c = load_case('D_small')
    ops = {'shrink_zone': 0.25, 'grow_zone': 0.25, 'swap_zone': 0.2}
    ops.update({'gibbs_sample_weights': 0.1, 'gibbs_sample_p_zones': 0.2} if sample_source else
               {'alter_weights': 0.1, 'alter_p_zones': 0.2})
    monkeypatch.setenv('SBAYES_B200_ZONE_CACHE', '0')
    ref, ll_ref, pr_ref = _run_chains(c, 2, 0, 400, ops, sample_source, team=1)
    monkeypatch.delenv('SBAYES_B200_ZONE_CACHE')
    for team in (1, 2):
        got, ll, pr = _run_chains(c, 2, 0, 400, ops, sample_source, team=team)
        assert np.allclose(ll, ll_ref, rtol=1e-10) and np.allclose(pr, pr_ref, rtol=1e-10, atol=1e-10), team
        for a, b in zip(ref, got):
            assert np.array_equal(a['zones'], b['zones']), team
            if sample_source:
                assert np.array_equal(a['source'], b['source']), team
    assert not np.array_equal(ref[0]['zones'], c['g']['zones'])            # the zones did move


@pytest.mark.parametrize('sample_source', [True, False])
def test_large_site_count_keeps_running_values_exact(sample_source, built_lib):
    """20000 sites (configs[4]'s largest shape): the adjacency CSR no longer fits in shared memory (the zone cache
    does), teams of 8 CTAs per chain, 1250 groups in the likelihood tables.  Running ll / prior / counts == a
    from-scratch evaluation after a zone-move-heavy mix, zones stay disjoint and within their size bounds."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    n, f, s_, z = 20000, 32, 4, 4
    data = synthetic.make_data(n, f, s_, n_families=3, seed=5, with_dist=False, planted_zones=z, planted_size=15)
    eng = Engine(state_index(data.features), data.states, data.families, data.network['adj_mat'], None, n_zones=z,
                 n_chains=3, inheritance=True, sample_source=sample_source)
    for ch in range(3):
        st = synthetic.random_state(data, z, 8, True, sample_source, seed=20 + ch)
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'],
                      np.argmax(st['source'], -1) if sample_source else None)
    eng.eval()
    ops = {'shrink_zone': 0.2, 'grow_zone': 0.3, 'swap_zone': 0.2}
    ops.update({'gibbs_sample_weights': 0.1, 'gibbs_sample_p_zones': 0.1, 'gibbs_sample_p_global': 0.05,
                'gibbs_sample_p_families': 0.05} if sample_source else
               {'alter_weights': 0.1, 'alter_p_zones': 0.1, 'alter_p_global': 0.05, 'alter_p_families': 0.05})
    eng.set_operators(ops, PRECISION, min_size=3, max_size=40, p_grow_connected=0.85, seed=9)
    eng.run(300)
    ll_run, pr_run = eng.chain_scalars()
    cnt_run = [eng.counts(ch, recount=False) for ch in range(3)] if sample_source else None
    ll_new, pr_new, _ = eng.eval()
    assert np.allclose(ll_run, ll_new, rtol=1e-9) and np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9)
    acc, _ = eng.op_stats()
    assert acc[:, :3].sum() > 0                                            # zone moves were accepted
    for ch in range(3):
        zz = eng.get_state(ch, fields=('zones',))['zones']
        assert (zz.sum(axis=0) <= 1).all() and (zz.sum(axis=1) >= 3).all() and (zz.sum(axis=1) <= 40).all()
        if sample_source:
            new = eng.counts(ch, recount=True)
            for k in new:
                if new[k] is not None:
                    assert np.array_equal(cnt_run[ch][k], new[k]), (ch, k)
    eng.close()


ODD_SHAPES = {
    # name: (N, F, S, Fam, Z, zone_size, ragged, max_size)
    'tiny': (12, 3, 2, 0, 1, 3, False, 12),                  # max_size = N, one zone, two weights (C = 2)
    'max_states': (150, 20, 32, 2, 2, 5, True, 30),          # S = SB_MAX_S, ragged state counts
    'many_families': (400, 24, 8, 120, 3, 6, False, 30),     # the sweep's tables exceed shared memory: fallbacks
    'one_feature_block': (200, 16, 5, 4, 7, 4, True, 20),    # F = 16: a single feature block, teams stay at 1
    'hundreds_of_families': (1500, 20, 6, 700, 2, 5, False, 20),   # no shared-memory table fits: every fallback kernel
    'wide': (120, 530, 3, 3, 2, 6, False, 25),               # F > 256 (several correction passes in K1), F % 16 != 0
}


@pytest.mark.parametrize('shape', sorted(ODD_SHAPES))
@pytest.mark.parametrize('sample_source', [True, False])
def test_odd_shapes_keep_running_values_exact(shape, sample_source, built_lib):
    """Corners of the shape space through the whole sampler: running ll / prior / count tables == from-scratch values,
    zones disjoint and within bounds, every launch succeeds (shared-memory layouts, fallbacks, team sizes)."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    n, f, s_, fam, z, zs, ragged, max_size = ODD_SHAPES[shape]
    inh = fam > 0
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=3, with_dist=False, ragged_states=ragged, na_frac=0.05)
    n_chains = 3
    eng = Engine(state_index(data.features), data.states, data.families if inh else None, data.network['adj_mat'], None,
                 n_zones=z, n_chains=n_chains, inheritance=inh, sample_source=sample_source)
    for ch in range(n_chains):
        st = synthetic.random_state(data, z, zs, inh, sample_source, seed=40 + ch)
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'] if inh else None,
                      np.argmax(st['source'], -1) if sample_source else None)
    # the evaluation kernels against the oracle on the last uploaded state (chain n_chains - 1)
    import sbayes_oracle as orc
    fam_b = np.asarray(data.families, bool)
    args = (data.features, fam_b, inh, st['zones'], st['weights'], st['p_global'], st['p_zones'],
            st['p_families'] if inh else None)
    ll_free = eng.eval(use_source=False)[0][n_chains - 1]
    ref_free = orc.log_likelihood(*args, source=None)
    assert abs(ll_free - ref_free) <= 1e-9 * abs(ref_free), (ll_free, ref_free)
    if sample_source:
        ll_src = eng.eval(use_source=True)[0][n_chains - 1]
        ref_src = orc.log_likelihood(*args, source=st['source'])
        assert abs(ll_src - ref_src) <= 1e-9 * abs(ref_src), (ll_src, ref_src)
    eng.eval()
    ops = {'shrink_zone': 0.15, 'grow_zone': 0.2, 'swap_zone': 0.15}
    gib = {'gibbs_sample_weights': 0.15, 'gibbs_sample_p_zones': 0.15, 'gibbs_sample_p_global': 0.1,
           'gibbs_sample_p_families': 0.05, 'gibbs_sample_sources': 0.05}
    alt = {'alter_weights': 0.15, 'alter_p_zones': 0.15, 'alter_p_global': 0.1, 'alter_p_families': 0.1}
    ops.update({k: v for k, v in (gib if sample_source else alt).items() if inh or 'families' not in k})
    eng.set_operators(ops, PRECISION, min_size=3, max_size=max_size, p_grow_connected=0.85, seed=13)
    for _ in range(2):
        eng.run(250)
        ll_run, pr_run = eng.chain_scalars()
        cnt_run = [eng.counts(ch, recount=False) for ch in range(n_chains)] if sample_source else None
        ll_new, pr_new, _ = eng.eval()
        assert np.all(np.isfinite(ll_new)), ll_new
        assert np.allclose(ll_run, ll_new, rtol=1e-9), (ll_run, ll_new)
        assert np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9), (pr_run, pr_new)
        for ch in range(n_chains):
            zz = eng.get_state(ch, fields=('zones',))['zones']
            assert (zz.sum(axis=0) <= 1).all() and (zz.sum(axis=1) >= 3).all() and (zz.sum(axis=1) <= max_size).all()
            if sample_source:
                new = eng.counts(ch, recount=True)
                for k in new:
                    if new[k] is not None:
                        assert np.array_equal(cnt_run[ch][k], new[k]), (ch, k)
    eng.close()


def test_more_chains_than_sms(built_lib):
    """200 chains on 148 SMs: one CTA per chain, the launch runs in two waves of persistent CTAs; every chain keeps
    its running values consistent and chains with different global ids differ."""
    c = load_case('B_balkan_like')
    n_chains = 200
    eng = make_engine(c, n_chains=n_chains, sample_source=True)
    for ch in range(n_chains):
        set_golden_state(eng, c, ch)
    eng.eval()
    eng.set_operators(MIX, PRECISION, min_size=3, max_size=20, p_grow_connected=0.85, seed=5)
    eng.run(200)
    ll_run, pr_run = eng.chain_scalars()
    ll_new, pr_new, _ = eng.eval()
    assert np.allclose(ll_run, ll_new, rtol=1e-9) and np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9)
    assert len(set(np.round(ll_run, 6))) > n_chains // 2
    for ch in (0, 147, 148, 199):
        a, b = eng.counts(ch, recount=False), eng.counts(ch, recount=True)
        for k in a:
            if a[k] is not None:
                assert np.array_equal(a[k], b[k]), (ch, k)
    eng.close()


@pytest.mark.parametrize('shape', ['tiny', 'many_families', 'wide'])
@pytest.mark.parametrize('sample_source', [True, False])
def test_warmup_mode_on_odd_shapes(shape, sample_source, built_lib):
    """Warm-up variant of the operators (zone_sampling.py:1272-1577: per-chain max_size / p_grow_connected, forced
    accepts of the source moves, quirk Q3) on corner shapes: every accepted zone move pays the source sweep and the
    zone-cache update, so this is the densest exercise of both; running values == from-scratch values."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    n, f, s_, fam, z, zs, ragged, max_size = ODD_SHAPES[shape]
    inh = fam > 0
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=3, with_dist=False, ragged_states=ragged, na_frac=0.05)
    n_chains = 4
    eng = Engine(state_index(data.features), data.states, data.families if inh else None, data.network['adj_mat'], None,
                 n_zones=z, n_chains=n_chains, inheritance=inh, sample_source=sample_source, warmup=True)
    for ch in range(n_chains):
        st = synthetic.random_state(data, z, zs, inh, sample_source, seed=60 + ch)
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'] if inh else None,
                      np.argmax(st['source'], -1) if sample_source else None)
    eng.eval()
    ops = {'shrink_zone': 0.2, 'grow_zone': 0.25, 'swap_zone': 0.15}
    ops.update({'gibbs_sample_weights': 0.2, 'gibbs_sample_p_zones': 0.2} if sample_source else
               {'alter_weights': 0.2, 'alter_p_zones': 0.2})
    sizes = np.linspace(max(4, (zs + max_size) // 4), max_size, n_chains).astype(np.int32)       # util.py:1182-1199
    eng.set_operators(ops, PRECISION, min_size=3, max_size=sizes, p_grow_connected=[0.95, 0.85, 0.95, 0.85], seed=21)
    eng.run(300)
    ll_run, pr_run = eng.chain_scalars()
    cnt_run = [eng.counts(ch, recount=False) for ch in range(n_chains)] if sample_source else None
    ll_new, pr_new, _ = eng.eval()
    assert np.all(np.isfinite(ll_new))
    assert np.allclose(ll_run, ll_new, rtol=1e-9) and np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9)
    acc, rej = eng.op_stats()
    assert acc[:, :3].sum() > 0
    for ch in range(n_chains):
        zz = eng.get_state(ch, fields=('zones',))['zones']
        assert (zz.sum(axis=0) <= 1).all() and (zz.sum(axis=1) >= 3).all()
        assert (zz.sum(axis=1) <= max(int(sizes[ch]), zs)).all(), (ch, zz.sum(axis=1), sizes)
        if sample_source:
            new = eng.counts(ch, recount=True)
            for k in new:
                if new[k] is not None:
                    assert np.array_equal(cnt_run[ch][k], new[k]), (ch, k)
    eng.close()


@pytest.mark.parametrize('shape', ['tiny', 'max_states', 'one_feature_block'])
@pytest.mark.parametrize('sample_source', [True, False])
def test_informative_priors_on_odd_shapes(shape, sample_source, built_lib):
    """Cost-based geo prior (Prim MST over the last zone, K3), uniform size prior and count-based Dirichlet priors on
    corner shapes: the prior of the uploaded state matches the oracle, and after a run the incrementally maintained
    prior (size + geo + Dirichlet caches) equals a from-scratch evaluation."""
    import sbayes_oracle as orc
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    n, f, s_, fam, z, zs, ragged, max_size = ODD_SHAPES[shape]
    inh = fam > 0
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=3, with_dist=True, ragged_states=ragged, na_frac=0.05,
                               counts_prior=True)
    gibbs_g, alpha_g = orc.dirichlet_prior_tables(np.array(data.prior_universal['counts'], float), 10.)
    gibbs_f = alpha_f = None
    if inh:
        gibbs_f, alpha_f = orc.dirichlet_prior_tables(np.array(data.prior_inheritance['counts'], float), 10.)
    cost = data.geo_prior['cost_matrix']
    n_chains = 2
    eng = Engine(state_index(data.features), data.states, data.families if inh else None, data.network['adj_mat'], cost,
                 n_zones=z, n_chains=n_chains, inheritance=inh, sample_source=sample_source)
    eng.set_priors(size_type='uniform', geo_type='cost_based', geo_scale=120.0, p_global_type='counts',
                   p_families_type='counts' if inh else 'uniform', gibbs_counts_global=gibbs_g, alpha_global=alpha_g,
                   gibbs_counts_fam=gibbs_f, alpha_fam=alpha_f)
    for ch in range(n_chains):
        st = synthetic.random_state(data, z, zs, inh, sample_source, seed=80 + ch)
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'] if inh else None,
                      np.argmax(st['source'], -1) if sample_source else None)
    _, prior, _ = eng.eval()
    cfg = {'geo': {'type': 'cost_based', 'scale': 120.0}, 'area_size': {'type': 'uniform'}, 'weights': {'type': 'uniform'},
           'contact': {'type': 'uniform'}, 'universal': {'type': 'counts'},
           'inheritance': {'type': 'counts' if inh else 'uniform'}}
    ref = orc.log_prior(st['zones'], st['p_global'], st['p_families'] if inh else None, data.states, inh, cfg,
                        cost_mat=cost, alpha_global=alpha_g, alpha_families=alpha_f)
    assert abs(prior[n_chains - 1] - ref) <= 1e-9 * max(1.0, abs(ref)), (prior, ref)
    ops = {'shrink_zone': 0.15, 'grow_zone': 0.2, 'swap_zone': 0.15}
    gib = {'gibbs_sample_weights': 0.1, 'gibbs_sample_p_zones': 0.1, 'gibbs_sample_p_global': 0.2,
           'gibbs_sample_p_families': 0.1}
    alt = {'alter_weights': 0.1, 'alter_p_zones': 0.1, 'alter_p_global': 0.2, 'alter_p_families': 0.1}
    ops.update({k: v for k, v in (gib if sample_source else alt).items() if inh or 'families' not in k})
    eng.set_operators(ops, PRECISION, min_size=3, max_size=max_size, p_grow_connected=0.85, seed=31)
    eng.run(400)
    ll_run, pr_run = eng.chain_scalars()
    ll_new, pr_new, _ = eng.eval()
    assert np.allclose(ll_run, ll_new, rtol=1e-9), (ll_run, ll_new)
    assert np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9), (pr_run, pr_new)
    s1 = eng.get_state(n_chains - 1, one_hot_source=False)
    ref = orc.log_prior(s1['zones'].astype(bool), s1['p_global'], s1['p_families'] if inh else None, data.states, inh, cfg,
                        cost_mat=cost, alpha_global=alpha_g, alpha_families=alpha_f)
    assert abs(pr_new[n_chains - 1] - ref) <= 1e-9 * max(1.0, abs(ref)), (pr_new, ref)
    eng.close()


@pytest.mark.parametrize('n_chains,inheritance,p_grow', [(1, True, 0.85), (2, True, 1.0), (3, False, 0.0)])
def test_bench_shape_with_few_chains(n_chains, inheritance, p_grow, built_lib):
    """The benchmark shape (2000 x 200 x 8, 7 zones, 20 families) with 1-3 chains: the automatic team size is 8 CTAs per
    chain (a whole GPC per chain); also without families (two mixture components) and with p_grow_connected at its
    extremes.  Running values == from-scratch values, count tables bit-exact, likelihood == oracle."""
    import sbayes_oracle as orc
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    data = synthetic.make_data(2000, 200, 8, n_families=20 if inheritance else 0, seed=0, with_dist=False,
                               planted_zones=7, planted_size=20)
    eng = Engine(state_index(data.features), data.states, data.families if inheritance else None,
                 data.network['adj_mat'], None, n_zones=7, n_chains=n_chains, inheritance=inheritance, sample_source=True)
    for ch in range(n_chains):
        st = synthetic.random_state(data, 7, 5, inheritance, True, seed=100 + ch)
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'],
                      st['p_families'] if inheritance else None, np.argmax(st['source'], -1))
    ops = {'shrink_zone': 0.1, 'grow_zone': 0.1, 'swap_zone': 0.05, 'gibbs_sample_weights': 0.3,
           'gibbs_sample_p_global': 0.1, 'gibbs_sample_p_zones': 0.25}
    if inheritance:
        ops['gibbs_sample_p_families'] = 0.1
    eng.eval()
    eng.set_operators(ops, PRECISION, min_size=3, max_size=50, p_grow_connected=p_grow, seed=3)
    eng.run(500)
    ll_run, pr_run = eng.chain_scalars()
    cnt_run = eng.counts(n_chains - 1, recount=False)
    ll_new, pr_new, _ = eng.eval()
    assert np.allclose(ll_run, ll_new, rtol=1e-9) and np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9)
    cnt_new = eng.counts(n_chains - 1, recount=True)
    for k in cnt_new:
        if cnt_new[k] is not None:
            assert np.array_equal(cnt_run[k], cnt_new[k]), k
    s1 = eng.get_state(n_chains - 1)
    ref = orc.log_likelihood(data.features, np.asarray(data.families, bool), inheritance, s1['zones'].astype(bool),
                             s1['weights'], s1['p_global'], s1['p_zones'], s1['p_families'] if inheritance else None,
                             source=s1['source'].astype(bool))
    assert abs(ll_new[n_chains - 1] - ref) <= 1e-9 * abs(ref), (ll_new, ref)
    acc, _ = eng.op_stats()
    assert acc[:, :3].sum() > 0
    eng.close()


@pytest.mark.parametrize('seed', range(int(os.environ.get('SB_FUZZ_SEEDS', '24'))))     # more seeds: SB_FUZZ_SEEDS=400
def test_random_shapes(seed, built_lib):
    """Seeded random shapes, priors and operator mixes (5..600 sites, 1..70 features, 2..12 states, 0..12 families,
    1..5 zones, both sampler modes, warm-up on / off, cost-based geo / size / count priors on / off): likelihood and
    prior of the uploaded state == oracle, running values == from-scratch values after a run."""
    import sbayes_oracle as orc
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine, state_index
    rng = np.random.default_rng(1000 + seed)
    z = int(rng.integers(1, 6))
    zs = int(rng.integers(3, 7))
    n = int(rng.integers(max(5, z * zs + 2), 600))
    f, s_ = int(rng.integers(1, 71)), int(rng.integers(2, 13))
    fam = int(rng.integers(0, 13)) if n > 30 else 0
    inh = fam > 0
    sample_source, warmup = bool(rng.integers(0, 2)), bool(rng.integers(0, 4) == 0)
    geo, size = bool(rng.integers(0, 2)), ['none', 'uniform', 'quadratic'][int(rng.integers(0, 3))]
    counts = bool(rng.integers(0, 2))
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=seed, with_dist=True, ragged_states=bool(rng.integers(0, 2)),
                               na_frac=float(rng.choice([0.0, 0.05, 0.3])), counts_prior=counts)
    cost = data.geo_prior['cost_matrix'] if geo else None
    n_chains = int(rng.integers(1, 5))
    try:
        states = [synthetic.random_state(data, z, zs, inh, sample_source, seed=seed * 10 + ch) for ch in range(n_chains)]
    except ValueError:
        pytest.skip('could not place the random zones')
    eng = Engine(state_index(data.features), data.states, data.families if inh else None, data.network['adj_mat'], cost,
                 n_zones=z, n_chains=n_chains, inheritance=inh, sample_source=sample_source, warmup=warmup)
    gg = ag = gf = af = None
    if counts:
        gg, ag = orc.dirichlet_prior_tables(np.array(data.prior_universal['counts'], float), 10.)
        if inh:
            gf, af = orc.dirichlet_prior_tables(np.array(data.prior_inheritance['counts'], float), 10.)
    scale = float(rng.uniform(20, 300))
    eng.set_priors(size_type=size, geo_type='cost_based' if geo else 'uniform', geo_scale=scale,
                   p_global_type='counts' if counts else 'uniform', p_families_type='counts' if (counts and inh) else 'uniform',
                   gibbs_counts_global=gg, alpha_global=ag, gibbs_counts_fam=gf, alpha_fam=af)
    for ch, st in enumerate(states):
        eng.set_state(ch, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'] if inh else None,
                      np.argmax(st['source'], -1) if sample_source else None)
    st = states[-1]
    ll, prior, _ = eng.eval()
    args = (data.features, np.asarray(data.families, bool), inh, st['zones'], st['weights'], st['p_global'], st['p_zones'],
            st['p_families'] if inh else None)
    ref = orc.log_likelihood(*args, source=st['source'] if sample_source else None)
    assert abs(ll[-1] - ref) <= 1e-9 * max(1.0, abs(ref)), (ll, ref)
    cfg = {'geo': {'type': 'cost_based' if geo else 'uniform', 'scale': scale}, 'area_size': {'type': size},
           'weights': {'type': 'uniform'}, 'contact': {'type': 'uniform'},
           'universal': {'type': 'counts' if counts else 'uniform'},
           'inheritance': {'type': 'counts' if (counts and inh) else 'uniform'}}
    refp = orc.log_prior(st['zones'], st['p_global'], st['p_families'] if inh else None, data.states, inh, cfg,
                         cost_mat=cost, alpha_global=ag, alpha_families=af)
    assert abs(prior[-1] - refp) <= 1e-9 * max(1.0, abs(refp)), (prior, refp)
    names = ['shrink_zone', 'grow_zone', 'swap_zone'] + (
        ['gibbs_sample_weights', 'gibbs_sample_p_global', 'gibbs_sample_p_zones', 'gibbs_sample_sources'] +
        (['gibbs_sample_p_families'] if inh else []) if sample_source else
        ['alter_weights', 'alter_p_global', 'alter_p_zones'] + (['alter_p_families'] if inh else []))
    if f * 1 < 1 or (not sample_source and int(data.states.sum(1).min()) < 2):
        names = [k for k in names if not k.startswith('alter_p')]
    ops = {k: float(w) for k, w in zip(names, rng.dirichlet(np.ones(len(names)))) if w > 0.02}
    max_size = int(rng.integers(zs + 1, max(zs + 2, min(n, 60))))
    eng.set_operators(ops, PRECISION, min_size=3, max_size=max_size, p_grow_connected=float(rng.uniform(0, 1)), seed=seed)
    eng.run(200)
    ll_run, pr_run = eng.chain_scalars()
    cnt_run = [eng.counts(ch, recount=False) for ch in range(n_chains)] if sample_source else None
    ll_new, pr_new, _ = eng.eval()
    ok = np.isfinite(ll_new)
    assert np.allclose(ll_run[ok], ll_new[ok], rtol=1e-9) and np.array_equal(np.isfinite(ll_run), ok), (ll_run, ll_new)
    assert np.allclose(pr_run, pr_new, rtol=1e-9, atol=1e-9), (pr_run, pr_new)
    for ch in range(n_chains):
        zz = eng.get_state(ch, fields=('zones',))['zones']
        assert (zz.sum(axis=0) <= 1).all() and (zz.sum(axis=1) >= 3).all() and (zz.sum(axis=1) <= max(max_size, zs)).all()
        if sample_source:
            new = eng.counts(ch, recount=True)
            for k in new:
                if new[k] is not None:
                    assert np.array_equal(cnt_run[ch][k], new[k]), (ch, k)
    eng.close()
