"""TEST INFRASTRUCTURE ONLY -- operator-level golden vectors from the UNMODIFIED reference:

  pair      dirichlet_proposal (zone_sampling.py:538-569) with `np.random.dirichlet` scripted to return a FORCED new
            pair: (c0, c1, n0, n1, precision, log_q, log_q_back) -- the parity surface of `sb_pair_log_q`
  warm_*    the warm-up variants of grow / shrink / swap (ZoneMCMCWarmup, :1328-1574) on the fixed states of the
            golden cases: (kind, zone, site_new, site_removed, rho, log_q, log_q_back) with resample_source=False
            (quirk Q3: shrink's q_back = 1 / (size + 1)), and the log_q the same operators return WITH source
            resampling (-inf = Q_GIBBS: warm-up zone moves are always accepted, :1293-1296)

Run in the build container:   python oracle/make_golden_ops.py
"""
import os
import random
import sys
from types import SimpleNamespace

import numpy as np
from scipy.sparse import csr_matrix

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import refshim  # noqa: E402
import sbayes_oracle as orc  # noqa: E402

ref = refshim.modules()
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
WARM_CASES = ['A_sim_like', 'B_balkan_like', 'C_samerica_like', 'F_balkan_data']


def pair_goldens(n=300, seed=5):
    rng = np.random.default_rng(seed)
    rows = []
    orig = np.random.dirichlet
    try:
        for _ in range(n):
            mass = float(rng.choice([1.0, 1.0, rng.uniform(0.02, 1.0)]))     # alter_p_*: the pair holds part of the row
            t0 = float(np.clip(rng.beta(0.7, 0.7), 1e-6, 1 - 1e-6))
            v0 = float(np.clip(rng.beta(0.7, 0.7), 1e-6, 1 - 1e-6))
            kappa = float(rng.choice([10, 15, 20, 30, 40]))
            w = np.array([t0, 1.0 - t0])
            w_new = np.array([v0, 1.0 - v0])
            np.random.dirichlet = lambda alpha, _w=w_new: _w.copy()
            got, lq, lqb = ref.zone_sampling.ZoneMCMC.dirichlet_proposal(w, kappa)
            assert np.array_equal(got, w_new)
            rows.append((w[0] * mass, w[1] * mass, w_new[0] * mass, w_new[1] * mass, kappa, lq, lqb))
    finally:
        np.random.dirichlet = orig
    return np.array(rows)


def load_data(name):
    g = dict(np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False))
    n, f = g['features_idx'].shape
    s = int(g['n_states'])
    inh = bool(g['inheritance'])
    adj = csr_matrix((np.ones(len(g['adj_indices']), dtype=np.int64), g['adj_indices'], g['adj_indptr']), shape=(n, n))
    cost = g['cost'] if g['cost'].size else None
    data = SimpleNamespace(features=orc.one_hot(g['features_idx'], s), families=g['families'].astype(bool),
                           states=g['states'].astype(bool), network={'adj_mat': adj, 'locations': None, 'dist_mat': cost},
                           prior_universal={}, prior_inheritance={}, geo_prior={'cost_matrix': cost}, is_simulated=False)
    # the warm-up proposals do not depend on the priors: uniform ones keep the model construction simple
    pcfg = {'geo': {'type': 'uniform'}, 'area_size': {'type': 'none'}, 'weights': {'type': 'uniform'},
            'universal': {'type': 'uniform'}, 'contact': {'type': 'uniform'}}
    if inh:
        pcfg['inheritance'] = {'type': 'uniform'}
    cfg = {'N_AREAS': g['zones'].shape[0], 'MIN_M': 3, 'MAX_M': 50, 'INHERITANCE': inh, 'SAMPLE_SOURCE': True, 'PRIOR': pcfg}
    return g, data, cfg, inh


def warm_goldens(name, seed=11):
    g, data, cfg, inh = load_data(name)
    model = ref.model.Model(data, cfg)
    sampler = ref.zone_sampling.ZoneMCMCWarmup(
        data=data, model=model, n_chains=2, operators={'grow_zone': 1.},
        var_proposal={'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20},
        p_grow_connected=0.85, initial_size=5)
    sampler.p_grow_connected = [0.85, 0.95]
    sampler.max_size = [50, 50]
    C = 3 if inh else 2
    src = np.eye(C, dtype=bool)[g['source_idx']]

    def sample():
        return ref.zone_sampling.Sample(zones=g['zones'].copy(), weights=g['weights'].copy(), p_global=g['p_global'].copy(),
                                        p_zones=g['p_zones'].copy(), p_families=g['p_families'].copy() if inh else None,
                                        source=src.copy())

    random.seed(seed)
    np.random.seed(seed)
    moves, src_logq = [], []
    for i, kind in enumerate(['grow', 'shrink', 'swap'] * 10):
        c = i % 2
        smp = sample()
        new, lq, lqb = getattr(sampler, kind + '_zone')(smp, c=c, resample_source=False)
        if new is smp:
            continue
        diff = new.zones.astype(int) - smp.zones.astype(int)
        zz = int(np.flatnonzero(np.abs(diff).sum(1))[0])
        added, removed = np.flatnonzero(diff[zz] == 1), np.flatnonzero(diff[zz] == -1)
        moves.append((['grow', 'shrink', 'swap'].index(kind), zz, int(added[0]) if len(added) else -1,
                      int(removed[0]) if len(removed) else -1, sampler.p_grow_connected[c], lq, lqb))
        smp = sample()
        new, lq_s, lqb_s = getattr(sampler, kind + '_zone')(smp, c=c, resample_source=True)
        if new is not smp:
            src_logq.append((lq_s, lqb_s))
    return np.array(moves, float), np.array(src_logq, float)


def main():
    save = {'pair': pair_goldens()}
    for name in WARM_CASES:
        mv, sq = warm_goldens(name)
        save['warm_' + name] = mv
        save['warm_src_' + name] = sq
        print(name, len(mv), 'warm-up moves; with source resampling log_q =', np.unique(sq[:, 0]), 'log_q_back =', np.unique(sq[:, 1]))
    out = os.path.join(GOLDEN, 'ops.npz')
    np.savez_compressed(out, **save)
    print('wrote', out, os.path.getsize(out), 'B; pair rows', save['pair'].shape)


if __name__ == '__main__':
    main()
