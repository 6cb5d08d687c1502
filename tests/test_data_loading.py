"""CPU: the input formats in front of the hot path (contact_zones_b200/data.py, experiment_setup.py) -- CSV features ->
uint8 state index / one-hot, families, Delaunay network, count priors, cost matrix, config defaults.  A synthetic
dataset is written to a temp dir (always runs); when the reference tree is present (this container, not the GPU box)
the same loaders are compared with the reference's own on its bundled Balkan data."""
import json
import os

import numpy as np
import pandas as pd
import pytest

from contact_zones_b200 import data as D
from contact_zones_b200.experiment_setup import Experiment

REF = '/root/reference'


def write_dataset(tmp, n=40, seed=0):
    rng = np.random.default_rng(seed)
    states = {'f_bin': ['no', 'yes'], 'f_tri': ['a', 'b', 'c'], 'f_col': ['red', 'green'], 'f_four': ['w', 'x', 'y', 'z']}
    fs = pd.DataFrame({k: v + [None] * (4 - len(v)) for k, v in states.items()})
    fs.to_csv(tmp / 'feature_states.csv', index=False)
    rows = []
    fams = ['Alpha', 'Beta', None]
    for i in range(n):
        r = {'id': f'L{i}', 'name': f' lang {i} ', 'x': rng.uniform(0, 100), 'y': rng.uniform(0, 100),
             'family': fams[i % 3]}
        for k, v in states.items():
            r[k] = None if rng.random() < 0.1 else ' ' + str(rng.choice(v)) + ' '       # padded strings get stripped
        rows.append(r)
    df = pd.DataFrame(rows)
    df.to_csv(tmp / 'features.csv', index=False)
    uni = pd.DataFrame({'feature': list(states), 'no': [30, 0, 0, 0], 'yes': [10, 0, 0, 0], 'a': [0, 5, 0, 0],
                        'b': [0, 6, 0, 0], 'c': [0, 7, 0, 0], 'red': [0, 0, 2, 0], 'green': [0, 0, 3, 0],
                        'w': [0, 0, 0, 1], 'x': [0, 0, 0, 1], 'y': [0, 0, 0, 1], 'z': [0, 0, 0, 50]})
    uni.to_csv(tmp / 'universal.csv', index=False)
    uni.assign(no=[3, 0, 0, 0]).to_csv(tmp / 'alpha_counts.csv', index=False)
    cost = pd.DataFrame(rng.uniform(1, 9, (n, n)), index=[f'L{i}' for i in range(n)][::-1],
                        columns=[f'L{i}' for i in range(n)][::-1])
    cost.to_csv(tmp / 'cost.csv')
    cfg = {'mcmc': {'N_STEPS': 400, 'N_SAMPLES': 4, 'WARM_UP': {'N_WARM_UP_STEPS': 50, 'N_WARM_UP_CHAINS': 3},
                    'STEPS': {'area': 0.1, 'weights': 0.3, 'universal': 0.1, 'contact': 0.3, 'inheritance': 0.2}},
           'model': {'N_AREAS': 2, 'INHERITANCE': True, 'MAX_M': 12,
                     'PRIOR': {'geo': {'type': 'cost_based', 'scale': 4.0, 'file': 'cost.csv'},
                               'universal': {'type': 'counts', 'file_type': 'counts_file', 'file': 'universal.csv',
                                             'scale_counts': 10},
                               'inheritance': {'type': 'counts', 'file_type': 'counts_file', 'scale_counts': 10,
                                               'files': {'Alpha': 'alpha_counts.csv'}}}},
           'data': {'FEATURES': 'features.csv', 'FEATURE_STATES': 'feature_states.csv'},
           'results': {'RESULTS_PATH': 'results'}}
    json.dump(cfg, open(tmp / 'config.json', 'w'))
    return df, states, cost


def test_loader_on_synthetic_csv(tmp_path):
    df, states, cost = write_dataset(tmp_path)
    exp = Experiment('unit', tmp_path / 'config.json', log=False)
    cfg = exp.config
    assert cfg['mcmc']['P_GROW_CONNECTED'] == 0.85 and cfg['model']['MIN_M'] == 3           # defaults filled in
    assert abs(sum(cfg['mcmc']['STEPS'].values()) - 1) < 1e-12 and cfg['mcmc']['STEPS']['source'] == 0.0
    assert cfg['mcmc']['N_CHAINS'] == 1 and cfg['model']['PRIOR']['area_size'] == {'type': 'none'}
    assert os.path.isabs(str(cfg['data']['FEATURES'])) and exp.path_results.is_dir()
    d = D.Data(exp)
    d.load_features(); d.load_universal_counts(); d.load_inheritance_counts(); d.load_geo_cost_matrix()
    n, f, s = 40, 4, 4
    assert d.features.shape == (n, f, s) and d.features.dtype == bool and d.features_idx.dtype == np.uint8
    assert np.array_equal(d.states, [[1, 1, 0, 0], [1, 1, 1, 0], [1, 1, 0, 0], [1, 1, 1, 1]])
    for j, (k, v) in enumerate(states.items()):
        want = np.array([255 if not isinstance(c, str) else v.index(c.strip()) for c in df[k].tolist()])
        assert np.array_equal(d.features_idx[:, j], want)
    assert np.array_equal(d.features.sum(-1) == 0, d.features_idx == 255)                    # NA = all-zero one-hot
    assert np.array_equal(np.argmax(d.features, -1)[d.features_idx != 255], d.features_idx[d.features_idx != 255])
    assert d.family_names['external'] == ['Alpha', 'Beta'] and d.families.shape == (2, n)
    assert d.families[0, 0] == 1 and d.families[1, 1] == 1 and d.families[:, 2].sum() == 0
    adj = d.network['adj_mat']
    assert adj.shape == (n, n) and (adj != adj.T).nnz == 0 and adj.diagonal().sum() == 0
    loc = d.network['locations']
    assert np.allclose(d.network['dist_mat'][3, 7], np.hypot(*(loc[3] - loc[7])))
    assert np.array_equal(d.prior_universal['counts'][0], [30, 10, 0, 0]) and d.prior_universal['counts'][3, 3] == 50
    pi = d.prior_inheritance['counts']
    assert pi.shape == (2, f, s) and pi[0, 0, 0] == 3 and pi[1].sum() == 0                   # no file -> zeros
    c = d.geo_prior['cost_matrix']                       # re-ordered to the site order and symmetrised
    raw = cost.loc[[f'L{i}' for i in range(n)], [f'L{i}' for i in range(n)]].to_numpy()
    assert np.allclose(c, (raw + raw.T) / 2)
    # the loaded object drives the model layer (constructors only: no device needed)
    from contact_zones_b200.model import Model
    m = Model(d, cfg['model'])
    g = m.prior.prior_p_global.counts
    assert np.allclose(g[0], [1 + 7.5, 1 + 2.5, 1, 1]) and np.allclose(g[3, 3], 1 + 50 * 10 / 53)    # scaled to 10
    kw = m.prior.device_config()
    assert kw['geo_type'] == 'cost_based' and kw['p_global_type'] == 'counts' and kw['alpha_fam'].shape == (2, f, s)


def test_config_errors(tmp_path):
    write_dataset(tmp_path)
    cfg = json.load(open(tmp_path / 'config.json'))
    bad = dict(cfg, model={k: v for k, v in cfg['model'].items() if k != 'N_AREAS'})
    json.dump(bad, open(tmp_path / 'bad1.json', 'w'))
    with pytest.raises(NameError):
        Experiment('x', tmp_path / 'bad1.json', log=False)
    bad = json.loads(json.dumps(cfg)); bad['mcmc']['N_SAMPLES'] = 3
    json.dump(bad, open(tmp_path / 'bad2.json', 'w'))
    with pytest.raises(ValueError):
        Experiment('x', tmp_path / 'bad2.json', log=False)
    bad = json.loads(json.dumps(cfg)); bad['simulation'] = {}
    json.dump(bad, open(tmp_path / 'bad3.json', 'w'))
    with pytest.raises(NotImplementedError):
        Experiment('x', tmp_path / 'bad3.json', log=False)


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, 'experiments', 'balkan')), reason='reference tree not present')
def test_loader_matches_reference_on_balkan():
    import refshim
    refshim.install()
    import sbayes.preprocessing as rp
    import sbayes.util as ru
    base = os.path.join(REF, 'experiments', 'balkan', 'data')
    ff, fs = os.path.join(base, 'features', 'features.csv'), os.path.join(base, 'features', 'feature_states.csv')
    ours = D.read_features_from_csv(ff, fs, return_index=True)
    theirs = ru.read_features_from_csv(ff, fs)
    assert np.array_equal(ours[0]['locations'], theirs[0]['locations'])
    assert np.array_equal(ours[2], theirs[2]) and np.array_equal(ours[5], theirs[5])          # features, states
    assert np.array_equal(ours[6], theirs[6]) and ours[7] == theirs[7]                         # families
    assert ours[4] == theirs[4] and list(ours[3]['external']) == list(theirs[3]['external'])
    assert ours[8] == theirs[8]                                                                # the log line
    net_o, net_t = D.compute_network(ours[0]), rp.compute_network(theirs[0])
    assert (net_o['adj_mat'] != net_t['adj_mat']).nnz == 0 and np.allclose(net_o['dist_mat'], net_t['dist_mat'])
    uc = os.path.join(base, 'prior_universal', 'universal_counts.csv')
    co, _ = D.read_universal_counts(ours[3], ours[4], uc, 'counts_file', fs)
    ct, _ = rp.read_universal_counts(theirs[3], theirs[4], uc, 'counts_file', fs)
    assert np.array_equal(co, ct)
    files = {'Greek': os.path.join(base, 'prior_inheritance', 'greek_counts.csv'),
             'Slavic': os.path.join(base, 'prior_inheritance', 'slavic_counts.csv')}
    io_, _ = D.read_inheritance_counts(ours[7], ours[3], ours[4], files, 'counts_file', fs)
    it, _ = rp.read_inheritance_counts(theirs[7], theirs[3], theirs[4], files, 'counts_file', fs)
    assert np.array_equal(io_, it)


def test_network_without_dense_distances():
    """compute_network(dense_distances=False): no N x N matrix; `Data.load_geo_cost_matrix` then hands the site
    coordinates to the geo prior (implicit Euclidean cost)."""
    from contact_zones_b200.data import compute_network
    rng = np.random.default_rng(1)
    loc = rng.random((40, 2)) * 100
    sites = {'id': list(range(40)), 'locations': loc, 'names': [f's{i}' for i in range(40)]}
    dense = compute_network(sites)
    lean = compute_network(sites, dense_distances=False)
    assert lean['dist_mat'] is None and (lean['adj_mat'] != dense['adj_mat']).nnz == 0
    assert np.array_equal(dense['dist_mat'], np.linalg.norm(loc[:, None] - loc, axis=-1))
