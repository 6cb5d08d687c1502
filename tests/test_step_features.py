"""GPU: features of the step kernel that are pinned on the reference or on the kernel's own invariants:

  * forced zone-move proposals run through `op_zone_move` itself (no twin kernel), zone cache on and off, and the
    WARM-UP variants against goldens from the reference's `ZoneMCMCWarmup` (quirk Q3, zone_sampling.py:1328-1574)
  * `sb_pair_log_q`: the pair-proposal code of alter_* against `dirichlet_proposal` (zone_sampling.py:538-569)
  * warm-up zone moves with SAMPLE_SOURCE are always accepted (the reference returns log_q = -inf, :1293-1296)
  * the device snapshot ring: one persistent launch with logging / swap events inside == one launch per event
  * sample_from_prior, per-chain error flags
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from helpers import PRECISION, load_case, make_engine, set_golden_state

pytestmark = pytest.mark.gpu
OPS = dict(np.load(os.path.join(GOLDEN, 'ops.npz')))
WARM_CASES = [k[5:] for k in OPS if k.startswith('warm_') and not k.startswith('warm_src_')]
KINDS = ['grow', 'shrink', 'swap']


@pytest.mark.parametrize('zone_cache', ['1', '0'])
@pytest.mark.parametrize('name', ['A_sim_like', 'B_balkan_like', 'C_samerica_like', 'F_balkan_data', 'G_samerica_data'])
def test_forced_moves_through_the_step_kernel(name, zone_cache, built_lib, monkeypatch):
    monkeypatch.setenv('SBAYES_B200_ZONE_CACHE', zone_cache)
    c = load_case(name)
    eng = make_engine(c, n_chains=3)
    for ch in range(3):
        set_golden_state(eng, c, ch)
    before = eng.get_state(1)
    for kind, z, sn, sr, lq, lqb in c['g']['ref_moves']:
        q, qb = eng.move_log_q(1, KINDS[int(kind)], int(z), int(sn), int(sr), 0.85)
        assert abs(q - lq) <= 1e-12 * max(1.0, abs(lq)) and abs(qb - lqb) <= 1e-12 * max(1.0, abs(lqb)), (kind, q, lq, qb, lqb)
    after = eng.get_state(1)
    for k in before:                                    # a forced proposal leaves the state alone
        assert (before[k] is None and after[k] is None) or np.array_equal(before[k], after[k])
    # the same answers once the sampler has been configured (another shared-memory layout / team size)
    ops = {'shrink_zone': .1, 'grow_zone': .1, 'swap_zone': .1, 'gibbs_sample_weights': .7}
    eng.set_operators(ops, PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=1)
    kind, z, sn, sr, lq, lqb = c['g']['ref_moves'][0]
    q, qb = eng.move_log_q(2, KINDS[int(kind)], int(z), int(sn), int(sr), 0.85)
    assert abs(q - lq) <= 1e-12 * max(1.0, abs(lq)) and abs(qb - lqb) <= 1e-12 * max(1.0, abs(lqb))
    eng.close()


@pytest.mark.parametrize('name', WARM_CASES)
def test_warmup_moves_equal_reference(name, built_lib):
    c = load_case(name)
    eng = make_engine(c, n_chains=1, warmup=True)
    set_golden_state(eng, c, 0)
    rows = OPS['warm_' + name]
    assert len(rows) >= 20 and {int(r[0]) for r in rows} == {0, 1, 2}
    for kind, z, sn, sr, rho, lq, lqb in rows:
        q, qb = eng.move_log_q(0, KINDS[int(kind)], int(z), int(sn), int(sr), float(rho))
        assert abs(q - lq) <= 1e-12 * max(1.0, abs(lq)) and abs(qb - lqb) <= 1e-12 * max(1.0, abs(lqb)), (kind, q, lq, qb, lqb)
    # quirk Q3 is in the goldens: the warm-up shrink answers 1 / (size + 1), not the grow-back probability
    shrink = rows[rows[:, 0] == 1][0]
    size = int(c['g']['zones'][int(shrink[1])].sum())
    assert abs(shrink[6] - np.log(1.0 / (size + 1))) < 1e-12
    # with source resampling the reference's warm-up operators return log_q = -inf (Q_GIBBS): always accepted
    assert np.all(np.isneginf(OPS['warm_src_' + name][:, 0])) and np.all(np.isfinite(OPS['warm_src_' + name][:, 1]))
    eng.close()


def test_warmup_zone_moves_are_always_accepted(built_lib):
    """ZoneMCMCWarmup with SAMPLE_SOURCE: every zone move that is not a forced reject (size limits, no candidate) is
    accepted (mcmc_generative.py:308-311 with log_q = -inf).  grow_zone alone therefore fills every zone to max_size."""
    c = load_case('C_samerica_like')
    eng = make_engine(c, n_chains=2, warmup=True)
    for ch in range(2):
        set_golden_state(eng, c, ch)
    eng.eval()
    max_size = 14
    eng.set_operators({'grow_zone': 1.0}, PRECISION, min_size=3, max_size=max_size, p_grow_connected=[0.85, 0.95], seed=3)
    size0 = c['g']['zones'].sum(axis=1)
    n = 400
    eng.run(n)
    acc, rej = eng.op_stats()
    for ch in range(2):
        sizes = eng.get_state(ch, fields=('zones',))['zones'].sum(axis=1)
        assert np.all(sizes == max_size)
        assert acc[ch, 1] == int((max_size - size0).sum()) and acc[ch, 1] + rej[ch, 1] == n
    eng.close()


def test_pair_log_q_equals_dirichlet_proposal(built_lib):
    c = load_case('A_sim_like')
    eng = make_engine(c)
    rows = OPS['pair']
    worst = 0.0
    for c0, c1, n0, n1, kappa, lq, lqb in rows:
        got = eng.pair_log_q(c0, c1, n0, n1, kappa)
        worst = max(worst, abs(got - (lq - lqb)) / max(1.0, abs(lq), abs(lqb)))
    assert worst < 1e-11, worst
    eng.close()


def _run_api(c, n_chains, mc3, seed, n_steps=3000, n_samples=30, sample_source=True, ops=None, precision=None, **patch):
    from helpers import case_data, case_model_config
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import ZoneMCMC
    data = case_data(c)
    model = Model(data, case_model_config(c, sample_source=sample_source))
    kind = 'gibbs_sample' if sample_source else 'alter'
    ops_override = ops
    ops = {'shrink_zone': 0.04, 'grow_zone': 0.04, 'swap_zone': 0.02, f'{kind}_weights': 0.4, f'{kind}_p_global': 0.1,
           f'{kind}_p_zones': 0.3}
    if c['inheritance']:
        ops[f'{kind}_p_families'] = 0.1
    ops = ops_override if ops_override is not None else ops
    smp = ZoneMCMC(data=data, model=model, n_chains=n_chains, operators=ops, var_proposal=precision or PRECISION,
                   p_grow_connected=0.85, initial_size=5, mc3=mc3, swap_period=70, chain_swaps=2, seed=seed,
                   show_screen_log=False, **patch.pop('ctor', {}))
    for k, v in patch.items():
        setattr(smp, k, v)
    smp.generate_samples(n_steps, n_samples)
    st = smp.statistics
    idx = list(smp.chain_idx)
    smp.close()
    return st, idx


@pytest.mark.parametrize('mc3', [False, True])
def test_snapshot_ring_equals_one_launch_per_event(mc3, built_lib):
    """The logged samples do not depend on how the steps are cut into launches: everything in ONE persistent launch
    (events + snapshot ring), a ring that holds a single logged sample per launch, and one launch per event give the
    same `statistics` bit for bit (the RNG is keyed by the step number)."""
    c = load_case('C_samerica_like')
    n_chains = 4 if mc3 else 1
    ref, idx0 = _run_api(c, n_chains, mc3, seed=11)
    assert len(ref['sample_zones']) == 30 and ref['sample_id'] == list(range(30))
    for patch in ({'MAX_EVENTS_PER_LAUNCH': 1}, {'RING_BYTES_PER_LAUNCH': 1}, {'MAX_EVENTS_PER_LAUNCH': 7}):
        st, idx = _run_api(c, n_chains, mc3, seed=11, **patch)
        assert idx == idx0 and st['accepted_swaps'] == ref['accepted_swaps'] and st['n_swaps'] == ref['n_swaps']
        for k in ('sample_zones', 'sample_weights', 'sample_p_global', 'sample_p_zones', 'sample_p_families'):
            assert all(np.array_equal(a, b) for a, b in zip(st[k], ref[k])), (patch, k)
        assert st['sample_likelihood'] == ref['sample_likelihood'] and st['sample_prior'] == ref['sample_prior']
        assert dict(st['accept_operator']) == dict(ref['accept_operator'])
    if mc3:
        assert ref['n_swaps'] == (3000 // 70) * 2 and ref['accepted_swaps'] > 0


def test_sample_from_prior(built_lib):
    """mcmc_generative.py:115-116: with sample_from_prior the likelihood is identically 0 -- logged as 0, absent from
    every Metropolis ratio.  Source-free mode, uniform priors: the alter_p_global rows then follow the flat Dirichlet,
    whatever the data say (with the likelihood they are pulled to the data's state frequencies)."""
    c = load_case('A_sim_like')
    ops = {'alter_p_global': 0.9, 'alter_weights': 0.1}
    wide = dict(PRECISION, universal=2)          # wide pair proposals: the flat target mixes within the run
    st, _ = _run_api(c, 1, False, seed=5, n_steps=40000, n_samples=400, sample_source=False, ops=ops, precision=wide,
                     ctor={'sample_from_prior': True})
    assert all(v == 0.0 for v in st['sample_likelihood'])
    pg = np.stack(st['sample_p_global'])[100:, 0]                     # (samples, F, S)
    napp = c['g']['states'].sum(axis=1)
    for f in range(c['F']):
        k = int(napp[f])
        # flat Dirichlet(1, ..., 1): mean 1 / k, standard deviation sqrt((k - 1) / (k^2 (k + 1)))
        assert np.all(np.abs(pg[:, f, :k].mean(axis=0) - 1.0 / k) < 0.08), (f, pg[:, f].mean(axis=0))
        sd = np.sqrt((k - 1) / (k * k * (k + 1.0)))
        assert np.all(np.abs(pg[:, f, :k].std(axis=0) - sd) < 0.35 * sd), (f, pg[:, f].std(axis=0), sd)
    # with the likelihood the same rows are held by the data of 120 sites: an order of magnitude narrower
    st2, _ = _run_api(c, 1, False, seed=5, n_steps=40000, n_samples=400, sample_source=False, ops=ops, precision=wide)
    assert all(v < 0.0 for v in st2['sample_likelihood'])
    assert np.stack(st2['sample_p_global'])[100:, 0].std(axis=0).max() < 0.1


def test_error_flags_reach_the_caller(built_lib):
    """A feature with ONE applicable state: the reference's alter_p_global raises ValueError in random.sample(states, 2)
    (zone_sampling.py:471); the device rejects the step and flags the chain, and the next scalar read fails."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200._lib import SB_ERR_NAPP, SbError
    from contact_zones_b200.engine import Engine
    data = synthetic.make_data(40, 3, 3, seed=1, with_dist=False)
    states = np.array(data.states, bool)
    states[1] = [True, False, False]
    feats = np.array(data.features, bool)
    feats[:, 1, :] = [True, False, False]
    eng = Engine(feats, states, None, data.network['adj_mat'], None, n_zones=1, n_chains=2, sample_source=False)
    st = synthetic.random_state(data, 1, 5, False, False, seed=2)
    pg = st['p_global'].copy(); pg[0, 1] = [1.0, 0.0, 0.0]
    pz = st['p_zones'].copy(); pz[0, 1] = [1.0, 0.0, 0.0]
    for ch in range(2):
        eng.set_state(ch, st['zones'], st['weights'], pg, pz)
    eng.eval()
    eng.set_operators({'alter_p_global': 1.0}, PRECISION, min_size=3, max_size=10, p_grow_connected=0.85, seed=1)
    eng.run(200)
    assert np.all(eng.chain_errors() == SB_ERR_NAPP)
    with pytest.raises(SbError, match='SB_E_NUMERIC'):
        eng.chain_scalars()
    eng.set_state(0, st['zones'], st['weights'], pg, pz)              # a new state clears the chain's flags
    assert list(eng.chain_errors()) == [0, SB_ERR_NAPP]
    eng.close()


def test_implicit_cost_matrix_equals_dense(built_lib):
    """the cost_based geo prior without the dense N x N matrix.  With the site coordinates
    (`sb_upload_locations`) the device derives the Euclidean costs itself: same prior values as with the dense
    `dist_mat` the reference uses (simulation.py:155-156), same sampler trajectory."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine
    data = synthetic.make_data(300, 20, 4, n_families=3, seed=9, with_dist=True, planted_zones=2, planted_size=10)
    loc, dense = data.network['locations'], data.geo_prior['cost_matrix']
    engs = [Engine(data.features, data.states, data.families, data.network['adj_mat'], dense, n_zones=2, n_chains=4,
                   inheritance=True, sample_source=True),
            Engine(data.features, data.states, data.families, data.network['adj_mat'], None, n_zones=2, n_chains=4,
                   inheritance=True, sample_source=True, locations=loc)]
    ops = {'shrink_zone': .15, 'grow_zone': .15, 'swap_zone': .1, 'gibbs_sample_weights': .2, 'gibbs_sample_p_global': .1,
           'gibbs_sample_p_zones': .2, 'gibbs_sample_p_families': .1}
    out = []
    for eng in engs:
        eng.set_priors(size_type='uniform', geo_type='cost_based', geo_scale=80.0)
        for c in range(4):
            st = synthetic.random_state(data, 2, 8, True, True, seed=40 + c)
            eng.set_state(c, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'], st['source'])
        ll, pr, geo = eng.eval()
        eng.set_operators(ops, PRECISION, min_size=3, max_size=25, p_grow_connected=0.85, seed=5)
        eng.run(4000)
        ll1, pr1 = eng.chain_scalars()
        ll2, pr2, geo2 = eng.eval()
        assert np.allclose(pr1, pr2, rtol=1e-9, atol=1e-9) and np.allclose(ll1, ll2, rtol=1e-9)
        out.append((pr, geo, pr1, geo2, [eng.get_state(c, fields=('zones',))['zones'] for c in range(4)]))
        eng.close()
    (pr_a, geo_a, pr1_a, geo2_a, z_a), (pr_b, geo_b, pr1_b, geo2_b, z_b) = out
    assert np.all(geo_a < 0) and np.allclose(geo_a, geo_b, rtol=1e-12, atol=0) and np.allclose(pr_a, pr_b, rtol=1e-12)
    assert all(np.array_equal(a, b) for a, b in zip(z_a, z_b))              # the same moves were accepted
    assert np.allclose(geo2_a, geo2_b, rtol=1e-12, atol=0)
    assert any(not np.array_equal(z, synthetic.random_state(data, 2, 8, True, True, seed=40 + c)['zones'])
               for c, z in enumerate(z_a))                                   # ... and zones did move


def test_cost_based_geo_prior_at_20000_sites(built_lib):
    """The dense cost matrix of 20000 sites would be 3.2 GB per GPU (and as much on the host); the implicit form needs
    320 KB.  Checked against scipy's MST over the last zone's members (oracle, model.py:1096-1139)."""
    import sbayes_oracle as orc
    from scipy.spatial.distance import cdist
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine
    data = synthetic.make_data(20000, 6, 3, n_families=0, seed=2, with_dist=False)
    loc = data.network['locations']
    eng = Engine(data.features, data.states, None, data.network['adj_mat'], None, n_zones=3, n_chains=2,
                 inheritance=False, sample_source=True, locations=loc)
    eng.set_priors(geo_type='cost_based', geo_scale=30.0)
    for c in range(2):
        st = synthetic.random_state(data, 3, 40, False, True, seed=70 + c)
        eng.set_state(c, st['zones'], st['weights'], st['p_global'], st['p_zones'], None, st['source'])
    ops = {'shrink_zone': .2, 'grow_zone': .2, 'swap_zone': .2, 'gibbs_sample_weights': .2, 'gibbs_sample_p_zones': .2}
    eng.eval()
    eng.set_operators(ops, PRECISION, min_size=3, max_size=120, p_grow_connected=0.85, seed=9)
    eng.run(600)
    _, pr, geo = eng.eval()
    for c in range(2):
        m = np.flatnonzero(eng.get_state(c, fields=('zones',))['zones'][-1])
        ref = orc.geo_prior_cost_based(np.ones((1, len(m)), bool), cdist(loc[m], loc[m]), 30.0)
        assert abs(geo[c] - ref) <= 1e-9 * abs(ref), (geo[c], ref)
    eng.close()
