"""Shared helpers of the test-suite: load a golden case and rebuild the reference-shaped arrays."""
import os

import numpy as np
from scipy.sparse import csr_matrix

import sbayes_oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load_case(name):
    g = dict(np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False))
    n, f = g['features_idx'].shape
    s = int(g['n_states'])
    c = {'name': name, 'g': g, 'N': n, 'F': f, 'S': s, 'inheritance': bool(g['inheritance'])}
    c['features'] = orc.one_hot(g['features_idx'], s)
    c['families'] = g['families'].astype(bool)
    c['adj'] = csr_matrix((np.ones(len(g['adj_indices']), dtype=np.int64), g['adj_indices'], g['adj_indptr']), shape=(n, n))
    c['cost'] = g['cost'] if g['cost'].size else None
    C = 3 if c['inheritance'] else 2
    c['C'] = C
    c['source'] = np.eye(C, dtype=bool)[g['source_idx']]
    c['p_families'] = g['p_families'] if c['inheritance'] else None
    c['Z'] = g['zones'].shape[0]
    c['geo_type'] = str(g['geo_type'])
    c['size_type'] = str(g['size_type'])
    c['geo_scale'] = float(g['geo_scale'])
    c['counts_prior'] = bool(g['counts_prior'])
    c['prior_cfg'] = {
        'geo': {'type': c['geo_type'], 'scale': c['geo_scale']}, 'area_size': {'type': c['size_type']},
        'weights': {'type': 'uniform'}, 'contact': {'type': 'uniform'},
        'universal': {'type': 'counts' if c['counts_prior'] else 'uniform'},
        'inheritance': {'type': 'counts' if c['counts_prior'] else 'uniform'}}
    return c


def rel_err(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(1e-300, np.maximum(np.abs(a), np.abs(b)))))


def make_engine(c, n_chains=1, sample_source=True, warmup=False, **kw):
    from contact_zones_b200.engine import Engine
    g = c['g']
    eng = Engine(g['features_idx'], g['states'], c['families'] if c['inheritance'] else None, c['adj'], c['cost'],
                 n_zones=c['Z'], n_chains=n_chains, inheritance=c['inheritance'], sample_source=sample_source,
                 warmup=warmup, **kw)
    eng.set_priors(size_type=c['size_type'], geo_type=c['geo_type'], geo_scale=c['geo_scale'],
                   p_global_type='counts' if c['counts_prior'] else 'uniform',
                   p_families_type='counts' if (c['counts_prior'] and c['inheritance']) else 'uniform',
                   gibbs_counts_global=g.get('gibbs_counts_global'), alpha_global=g.get('alpha_global'),
                   gibbs_counts_fam=g.get('gibbs_counts_fam'), alpha_fam=g.get('alpha_fam'))
    return eng


def set_golden_state(eng, c, chain=0, with_source=True):
    g = c['g']
    eng.set_state(chain, g['zones'], g['weights'], g['p_global'], g['p_zones'], c['p_families'],
                  g['source_idx'] if with_source else None)


# ---- posterior-summary fixtures of the reference sampler (oracle/make_golden_runs.py) -------------------------
PRECISION = {'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20}
RUN_KEYS = ['zone_size', 'planted_occ', 'w_comp', 'll_mean', 'prior_mean', 'acc']
# the per-site / per-feature summaries north_star names ("zone membership frequencies, weight means"): one comparison
# per site, per (feature, component), per (feature, state)
VECTOR_KEYS = ['occ_freq', 'w_mean', 'pg_mean']
# spread floors of the vector statistics: a site that no run of either side ever visits (or always holds) has zero
# spread on both sides; differences below 1 % occupancy / 0.004 in a mean weight or probability are not resolved by
# runs of this length
VECTOR_FLOOR = {'occ_freq': 1e-2, 'w_mean': 4e-3, 'pg_mean': 4e-3}


def run_cases():
    return sorted(f[:-4] for f in os.listdir(GOLDEN) if f.startswith('run_') and f.endswith('.npz'))


def load_run(name):
    g = dict(np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False))
    n, f = g['features_idx'].shape
    s = int(g['n_states'])
    c = {'name': name, 'g': g, 'N': n, 'F': f, 'S': s, 'inheritance': bool(g['inheritance']),
         'sample_source': bool(g['sample_source']), 'counts_prior': bool(g['counts_prior']),
         'geo_type': str(g['geo_type']), 'size_type': str(g['size_type']), 'geo_scale': float(g['geo_scale'])}
    c['features'] = orc.one_hot(g['features_idx'], s)
    c['families'] = g['families'].astype(bool)
    c['adj'] = csr_matrix((np.ones(len(g['adj_indices']), dtype=np.int64), g['adj_indices'], g['adj_indptr']), shape=(n, n))
    c['cost'] = g['cost'] if g['cost'].size else None
    if 'locations' in g and c['geo_type'] == 'cost_based':
        # the fixture keeps the site locations; the cost matrix is their Euclidean distance matrix
        # (simulation.py:155-156, checked bit for bit when the fixture was written)
        from scipy.spatial.distance import cdist
        c['cost'] = cdist(g['locations'], g['locations'])
    c['Z'] = g['zones'].shape[0]
    c['C'] = 3 if c['inheritance'] else 2
    c['p_families'] = g['p_families'] if c['inheritance'] else None
    c['operators'] = {str(k): float(v) for k, v in zip(g['op_names'], g['op_weights'])}
    c['precision'] = (dict(zip(('weights', 'universal', 'contact', 'inheritance'), (float(v) for v in g['precision'])))
                      if 'precision' in g else dict(PRECISION))
    c['prior_cfg'] = {
        'geo': {'type': c['geo_type'], 'scale': c['geo_scale']}, 'area_size': {'type': c['size_type']},
        'weights': {'type': 'uniform'}, 'contact': {'type': 'uniform'},
        'universal': {'type': 'counts' if c['counts_prior'] else 'uniform'},
        'inheritance': {'type': 'counts' if c['counts_prior'] else 'uniform'}}
    return c


def summarise_run(zones, weights, ll, prior, acc, planted, burn, p_global=None):
    """The label-switching-invariant summaries of oracle/make_golden_runs.py:summarise for one chain."""
    zones, weights = np.asarray(zones)[burn:], np.asarray(weights)[burn:]
    occ = zones.any(axis=1)
    out = {'zone_size': occ.sum(axis=1).mean(), 'planted_occ': occ[:, planted].mean(),
           'w_comp': weights.mean(axis=(0, 1)), 'll_mean': np.mean(np.asarray(ll)[burn:]),
           'prior_mean': np.mean(np.asarray(prior)[burn:]), 'acc': np.asarray(acc, dtype=float),
           'occ_freq': occ.mean(axis=0), 'w_mean': weights.mean(axis=0)}
    if p_global is not None:
        pg = np.asarray(p_global)[burn:]
        out['pg_mean'] = pg.mean(axis=0)[0] if pg.ndim == 4 else pg.mean(axis=0)
    return out


def vector_z_scores(c, chains, keys=None):
    """Per-element Welch z of the mean over `chains` against the reference runs' mean for the per-site / per-feature
    summaries, with the bar that belongs to the number of comparisons: `max|z|` must stay below the two-sided
    Student-t quantile at a family-wise 1e-3 (Bonferroni over the K elements, Welch-Satterthwaite degrees of freedom
    of the least favourable element), and the mean of z^2 -- which a bias confined to a few sites or features moves
    little but a diffuse bias moves a lot -- below 3.  Returns {key: (max|z|, bar, mean z^2)}."""
    from scipy import stats
    g = c['g']
    out = {}
    for k in (keys or VECTOR_KEYS):
        if 'ref_' + k not in g or k not in chains[0]:
            continue
        ref = np.asarray(g['ref_' + k], dtype=float)
        ours = np.stack([np.asarray(s[k], dtype=float) for s in chains])
        ref, ours = ref.reshape(len(ref), -1), ours.reshape(len(ours), -1)
        n1, n2 = len(ref), len(ours)
        floor = VECTOR_FLOOR[k]
        rs = np.maximum(ref.std(0, ddof=1), floor)
        os_ = np.maximum(ours.std(0, ddof=1), floor)
        a, b = rs ** 2 / n1, os_ ** 2 / n2
        z = np.abs(ours.mean(0) - ref.mean(0)) / np.sqrt(a + b)
        df = (a + b) ** 2 / (a ** 2 / (n1 - 1) + b ** 2 / (n2 - 1))
        bar = float(stats.t.ppf(1.0 - 1e-3 / (2 * z.size), df.min()))
        out[k] = (float(z.max()), bar, float(np.mean(z ** 2)))
    return out


def run_z_scores(c, chains):
    """z-score of the mean over `chains` (list of summarise_run dicts) against the reference runs' mean, using the
    run-to-run spread of both sides; the spread is floored (1e-3 + 0.5 % of the mean) so that a statistic on which all
    six reference runs happen to agree does not blow the score up."""
    g = c['g']
    out = {}
    for k in RUN_KEYS:
        ref = np.asarray(g['ref_' + k], dtype=float)
        ours = np.stack([np.asarray(s[k], dtype=float) for s in chains])
        rm, om = ref.mean(0), ours.mean(0)
        floor = 1e-3 + 5e-3 * np.abs(rm)
        rs = np.maximum(ref.std(0, ddof=1), floor)
        os_ = np.maximum(ours.std(0, ddof=1), floor) if len(chains) > 1 else rs
        out[k] = np.abs(om - rm) / np.sqrt(rs ** 2 / len(ref) + os_ ** 2 / len(chains))
    return out


# ---- reference-shaped `data` / config objects for the API mirror (contact_zones_b200.model / .sampling) ----------
def case_data(c):
    """types.SimpleNamespace with the attributes the hot path reads."""
    from types import SimpleNamespace
    g = c['g']
    f, s = c['F'], c['S']
    data = SimpleNamespace(
        features=c['features'], families=c['families'], states=g['states'].astype(bool),
        network={'adj_mat': c['adj'], 'locations': None, 'dist_mat': c['cost']},
        prior_universal={}, prior_inheritance={}, geo_prior={'cost_matrix': c['cost']},
        state_names={'internal': [list(range(int(k))) for k in g['states'].sum(1)],
                     'external': [[f's{j}' for j in range(int(k))] for k in g['states'].sum(1)]},
        feature_names={'external': [f'f{i}' for i in range(f)]},
        family_names={'external': [f'fam{i}' for i in range(c['families'].shape[0])]},
        is_simulated=False)
    if c['counts_prior']:
        data.prior_universal = {'counts': g['gibbs_counts_global'] - 1.}
        if c['inheritance']:
            data.prior_inheritance = {'counts': g['gibbs_counts_fam'] - 1.}
    return data


def case_model_config(c, sample_source=True, min_m=3, max_m=30):
    pc = {k: dict(v) for k, v in c['prior_cfg'].items()}
    for k in ('universal', 'inheritance'):
        if pc[k]['type'] == 'counts':
            # fixtures of the bundled data sets carry their config's scale_counts (NaN = null: counts used as they are)
            sc = float(c['g'].get('scale_' + k, 10.))
            pc[k].update(scale_counts=None if np.isnan(sc) else sc, file='golden', files={})
    return {'N_AREAS': c['Z'], 'MIN_M': min_m, 'MAX_M': max_m, 'INHERITANCE': c['inheritance'],
            'SAMPLE_SOURCE': sample_source, 'PRIOR': pc}
