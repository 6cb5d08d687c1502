"""Value / byte parity of the callers and file formats either side of the hot path against
fixtures written by the UNMODIFIED reference's `MCMC` facade (oracle/make_golden_results.py):

  CPU   `match_areas` + `rank_areas` reproduce the reference's relabelled lists (postprocessing.py:205-268, 316-365);
        `samples2file` writes stats_*.txt / areas_*.txt (and the ground-truth pair of a simulated data set) BYTE FOR
        BYTE like util.py:846-907 from the same `statistics` dict
  GPU   `contribution_per_area` (batched sb_eval over (sample, zone) pairs) equals postprocessing.py:271-313 at 1e-9,
        `MCMC.eval_ground_truth` equals mcmc_setup.py:122-172
"""
import json
import os
import pathlib
from types import SimpleNamespace

import numpy as np
import pytest
from scipy.sparse import csr_matrix
from scipy.spatial.distance import cdist

import sbayes_oracle as orc
from conftest import GOLDEN

CASES = ['results_balkan', 'results_simexp1']
RTOL = 1e-9


def load(name):
    g = dict(np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False))
    cfg = json.loads(str(g['config']))
    names = json.loads(str(g['names']))
    n, f = g['features_idx'].shape
    s = int(g['n_states'])
    inh = cfg['model']['INHERITANCE']
    loc = g['locations']
    dist = cdist(loc, loc)
    adj = csr_matrix((np.ones(len(g['adj_indices']), dtype=np.int64), g['adj_indices'], g['adj_indptr']), shape=(n, n))
    data = SimpleNamespace(
        features=orc.one_hot(g['features_idx'], s), states=g['states'].astype(bool), families=g['families'].astype(bool),
        network={'adj_mat': adj, 'locations': loc, 'dist_mat': dist}, geo_prior={'cost_matrix': dist},
        prior_universal={}, prior_inheritance={}, feature_names=names['feature_names'],
        state_names=names['state_names'], family_names=names['family_names'], is_simulated=bool(g['is_simulated']))
    pc = cfg['model']['PRIOR']
    if 'raw_counts_universal' in g:
        data.prior_universal = {'counts': g['raw_counts_universal'].copy()}
        pc['universal'].update(file='golden')
    if 'raw_counts_inheritance' in g:
        data.prior_inheritance = {'counts': g['raw_counts_inheritance'].copy()}
        pc['inheritance'].update(files={})
    if data.is_simulated:
        data.areas = g['true_zones']
        data.weights = g['data_weights']
        data.p_universal = g['true_p_global'][0]
        data.p_contact = g['true_p_zones']
        data.p_inheritance = g['true_p_families'] if 'true_p_families' in g else None
    return g, cfg, data


def statistics(g, tag, inh, single_from_ref=True):
    """The `statistics` dict of the reference run: 'pre' = as logged, 'post' = after match_areas + rank_areas."""
    n = len(g['likelihood'])
    st = {'sample_zones': [g[f'{tag}_zones'][i] for i in range(n)],
          'sample_p_zones': [g[f'{tag}_p_zones'][i] for i in range(n)],
          'sample_weights': [g['weights'][i] for i in range(n)], 'sample_p_global': [g['p_global'][i] for i in range(n)],
          'sample_p_families': [g['p_families'][i] if inh else None for i in range(n)],
          'sample_likelihood': [float(v) for v in g['likelihood']], 'sample_prior': [float(v) for v in g['prior']]}
    if single_from_ref:
        for k in ('lh', 'prior', 'posterior'):
            st[f'sample_{k}_single_zones'] = [[float(v) for v in row] for row in g[f'{tag}_{k}_single']]
    return st


@pytest.mark.parametrize('name', CASES)
def test_match_and_rank_equal_reference(name):
    from contact_zones_b200 import results
    g, cfg, _ = load(name)
    st = statistics(g, 'pre', cfg['model']['INHERITANCE'])
    st = results.rank_areas(results.match_areas(st))
    assert np.array_equal(np.stack(st['sample_zones']), g['post_zones'])
    assert np.array_equal(np.stack(st['sample_p_zones']), g['post_p_zones'])
    for k in ('lh', 'prior', 'posterior'):
        assert np.array_equal(np.asarray(st[f'sample_{k}_single_zones']), g[f'post_{k}_single'])


@pytest.mark.parametrize('name', CASES)
def test_samples2file_bytes_equal_reference(name, tmp_path):
    from contact_zones_b200 import results
    g, cfg, data = load(name)
    st = statistics(g, 'post', cfg['model']['INHERITANCE'])
    if data.is_simulated:
        st.update(true_zones=g['true_zones'], true_weights=g['true_weights'], true_p_global=g['true_p_global'],
                  true_p_zones=g['true_p_zones'], true_p_families=data.p_inheritance, true_ll=float(g['true_ll']),
                  true_prior=float(g['true_prior']), true_lh_single_zones=[float(v) for v in g['true_lh_single']],
                  true_prior_single_zones=[float(v) for v in g['true_prior_single']],
                  true_posterior_single_zones=[float(a) + float(b) for a, b in zip(g['true_lh_single'], g['true_prior_single'])])
    want = {str(k): g[f'file_{i}'].tobytes() for i, k in enumerate(g['file_names'])}
    key = {'stats_': 'parameters', 'areas_': 'areas'}
    paths = {}
    for rel in want:
        p = pathlib.Path(tmp_path) / rel
        p.parent.mkdir(parents=True, exist_ok=True)
        base = os.path.basename(rel)
        if 'ground_truth' in rel:
            paths['gt' if base.startswith('stats') else 'gt_areas'] = p
        else:
            paths[key[base[:6]]] = p
    results.samples2file(st, data, cfg, paths)
    for rel, ref_bytes in want.items():
        ours = (pathlib.Path(tmp_path) / rel).read_bytes()
        assert ours == ref_bytes, (rel, ours[:300], ref_bytes[:300])


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_contribution_per_area_equals_reference(name, built_lib):
    from contact_zones_b200 import results
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import Sample, ZoneMCMC
    g, cfg, data = load(name)
    inh = cfg['model']['INHERITANCE']
    model = Model(data, cfg['model'])
    sampler = SimpleNamespace(statistics=statistics(g, 'pre', inh, single_from_ref=False), model=model, data=data)
    results.contribution_per_area(sampler)
    for k in ('lh', 'prior', 'posterior'):
        got = np.asarray(sampler.statistics[f'sample_{k}_single_zones'])
        ref = g[f'pre_{k}_single']
        assert got.shape == ref.shape
        assert np.max(np.abs(got - ref) / np.maximum(1.0, np.abs(ref))) <= RTOL, k
    # the logged (ll, prior) of the reference run are the model's values of the logged states too
    for i in (0, len(g['prior']) - 1):
        s = Sample(g['pre_zones'][i], g['weights'][i], g['p_global'][i], g['pre_p_zones'][i],
                   g['p_families'][i] if inh else None)
        assert abs(model.prior(s) - g['prior'][i]) <= RTOL * max(1.0, abs(g['prior'][i]))
    if data.is_simulated:
        # MCMC.eval_ground_truth (mcmc_setup.py:122-172) through the facade's own method
        import logging
        from contact_zones_b200.mcmc_setup import MCMC
        exp = SimpleNamespace(config=cfg, path_results=pathlib.Path('/tmp'), logger=logging.getLogger('t'))
        mc = MCMC(data, exp)
        mc.samples = {}
        mc.sampler = ZoneMCMC(n_chains=1, data=data, model=mc.model, operators=mc.ops,
                              var_proposal=cfg['mcmc']['PROPOSAL_PRECISION'], p_grow_connected=0.85, initial_size=5)
        mc.eval_ground_truth()
        assert abs(mc.samples['true_ll'] - float(g['true_ll'])) <= RTOL * abs(float(g['true_ll']))
        assert abs(mc.samples['true_prior'] - float(g['true_prior'])) <= RTOL * max(1.0, abs(float(g['true_prior'])))
        assert np.allclose(mc.samples['true_lh_single_zones'], g['true_lh_single'], rtol=RTOL, atol=0)
        assert np.allclose(mc.samples['true_weights'], g['true_weights'], rtol=1e-15, atol=0)
