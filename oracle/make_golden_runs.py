"""TEST INFRASTRUCTURE ONLY -- posterior summaries of the UNMODIFIED reference sampler on small seeded
problems -> `tests/golden/run_*.npz`, the distributional parity fixtures (BASELINE.json north_star: "posterior
summaries ... must agree with the reference's CPU run within Monte-Carlo error").

Run in the build container (where /root/reference exists):   python oracle/make_golden_runs.py [case ...]
For every case the reference's `ZoneMCMC.generate_samples` (mcmc_generative.py:149) is run `N_RUNS` times
from the same initial state with different seeds of its two global generators; each run is reduced to
    occ_freq (N)      fraction of samples in which the site is in any zone   (label-switching invariant)
    zone_size         mean total number of sites in zones;  planted_occ  mean occupancy of the planted sites
    w_mean (F,C)      mean weights            pg_mean (F,S)   mean p_global
    ll_mean, prior_mean, acc (per operator acceptance ratio)
The fixtures keep every run's summary so that the tests can form a z-score against the run-to-run spread.
The same script runs `oracle/sbayes_sampler_oracle.py` on the same problem and prints both step rates.
"""
import multiprocessing as mp
import os
import random
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import refshim  # noqa: E402
import sbayes_oracle as orc  # noqa: E402
import sbayes_sampler_oracle as so  # noqa: E402
from contact_zones_b200 import synthetic  # noqa: E402

ref = refshim.modules()

PRECISION = {'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20}
N_RUNS = 6
# area-heavy operator mix so that the zone statistics mix within the run length (STEPS is an input of the path)
STEPS = {'area': 0.3, 'weights': 0.3, 'universal': 0.05, 'contact': 0.3, 'inheritance': 0.05, 'source': 0.0}

# name: dict(N, F, S, Fam, Z, inheritance, sample_source, counts_prior, geo, size, n_steps, n_samples)
CASES = {
    'run_A_source': dict(N=80, F=12, S=3, Fam=0, Z=1, inh=False, src=True, cp=False, geo='uniform', size='uniform',
                         n_steps=60000, n_samples=3000),
    'run_B_families': dict(N=60, F=12, S=3, Fam=3, Z=2, inh=True, src=True, cp=True, geo='uniform', size='none',
                           n_steps=60000, n_samples=3000),
    'run_C_alter': dict(N=60, F=10, S=4, Fam=3, Z=1, inh=True, src=False, cp=False, geo='cost_based',
                        size='quadratic', n_steps=60000, n_samples=3000),
    # the reference's BUNDLED data sets (BASELINE.json configs[1], configs[2]), read with the reference's own loaders
    # (oracle/refdata.py); count priors and scale_counts as in the experiment's config.json
    'run_D_balkan': dict(bundled='balkan', Z=2, inh=True, src=True, cp=True, geo='uniform', size='none',
                         n_steps=60000, n_samples=3000, max_m=10),      # 28 sites: MAX_M = 30 would let two zones take them all
    'run_E_samerica': dict(bundled='south_america', Z=3, inh=True, src=True, cp=True, geo='uniform', size='uniform',
                           n_steps=60000, n_samples=3000),
    # experiments/simulation/sim_exp1 (BASELINE.json configs[0]): its 951 bundled sites, features drawn by the
    # reference's own simulator (seeded; strength 1 = I_CONTACT 3 / E_CONTACT 0.5, area 4 of run_sim_exp1.py),
    # model as in its config.json (1 area, MAX_M 200, no families, uniform priors); starts at the simulated area
    'run_F_simexp1': dict(simulated='sim_exp1', Z=1, inh=False, src=True, cp=False, geo='uniform', size='none',
                          n_steps=40000, n_samples=2000, max_m=200),
    # ---- full-form fixtures (BASELINE.json configs[1], configs[2]): the bundled experiments with the number of areas,
    # MC3 chains, operator mix (STEPS), PROPOSAL_PRECISION, M_INITIAL and MAX_M of their own config.json, run through
    # ZoneMCMC(mc3=True, swap_period, chain_swaps) -- the logged chain is chain_idx[0] (mcmc_generative.py:214-223)
    'run_G_balkan_mc3': dict(bundled='balkan', Z=3, inh=True, src=True, cp=True, geo='uniform', size='none',
                             n_steps=60000, n_samples=3000, max_m=30, use_cfg=True, n_chains=4, mc3=(100, 3), n_runs=8),
    'run_H_samerica_mc3': dict(bundled='south_america', Z=5, inh=True, src=True, cp=True, geo='uniform', size='none',
                               n_steps=40000, n_samples=2000, max_m=30, use_cfg=True, n_chains=8, mc3=(100, 3),
                               n_runs=8, zone_size0=5),
    # experiments/simulation/sim_exp2 (its config.json): 951 bundled sites, 50 features and one family drawn by the
    # reference's seeded simulator, SAMPLE_SOURCE = false (alter_* operators), cost-based geo prior (scale 100),
    # 1 area, MAX_M 200, the config's STEPS / PROPOSAL_PRECISION; starts at the simulated area
    'run_I_simexp2': dict(simulated='sim_exp2', Z=1, inh=True, src=False, cp=False, geo='cost_based', size='none',
                          n_steps=40000, n_samples=2000, max_m=200, use_cfg=True, n_runs=8, skip_port=True,
                          cost_from_locations=True),
}


def prior_config(c):
    su, sf = c.get('scales', (10., 10.))
    cfg = {'geo': {'type': c['geo'], 'scale': 100.}, 'area_size': {'type': c['size']}, 'weights': {'type': 'uniform'},
           'universal': {'type': 'counts', 'scale_counts': su} if c['cp'] else {'type': 'uniform'},
           'contact': {'type': 'uniform'}}
    if c['inh']:
        cfg['inheritance'] = {'type': 'counts', 'scale_counts': sf} if c['cp'] else {'type': 'uniform'}
    return cfg


def summarise(zones, weights, p_global, ll, prior, burn, planted):
    zones, weights, p_global = np.asarray(zones)[burn:], np.asarray(weights)[burn:], np.asarray(p_global)[burn:]
    occ = zones.any(axis=1)
    return dict(occ_freq=occ.mean(axis=0), zone_size=occ.sum(axis=1).mean(), w_mean=weights.mean(axis=0),
                planted_occ=occ[:, planted].mean(), w_comp=weights.mean(axis=(0, 1)),
                pg_mean=p_global.mean(axis=0)[0] if p_global.ndim == 4 else p_global.mean(axis=0),
                ll_mean=np.mean(np.asarray(ll)[burn:]), prior_mean=np.mean(np.asarray(prior)[burn:]))


def build(case, seed=7):
    c = dict(CASES[case])
    if 'bundled' in c:
        import refdata
        data, info = refdata.load_bundled(c['bundled'])
        c.update(N=data.features.shape[0], F=data.features.shape[1], S=data.features.shape[2],
                 Fam=data.families.shape[0], scales=(info['scale_universal'], info['scale_inheritance']))
        # "planted" = the sites of the initial zones: planted_occ then tracks how much of the start survives
        data.planted_zones = synthetic.random_state(data, c['Z'], c.get('zone_size0', 5), c['inh'], False,
                                                    seed=seed + 1)['zones']
    elif 'simulated' in c:
        import refdata
        if c['simulated'] == 'sim_exp1':
            data, info = refdata.load_simulated(c['simulated'], seed=seed)
        else:
            data, info = refdata.load_simulated(c['simulated'], seed=seed, i_contact=None, e_contact=None, area=None,
                                                inheritance=c['inh'])
        c.update(N=data.features.shape[0], F=data.features.shape[1], S=data.features.shape[2],
                 Fam=data.families.shape[0])
    else:
        data = synthetic.make_data(c['N'], c['F'], c['S'], n_families=c['Fam'], na_frac=0.03, seed=seed,
                                   counts_prior=c['cp'], planted_zones=max(1, c['Z']), planted_size=10,
                                   planted_conc=0.3)
    c['steps'], c['precision'], c['m_initial'] = dict(STEPS), dict(PRECISION), 5
    if c.get('use_cfg'):
        m = info['mcmc']
        c['steps'] = dict(m['STEPS'], source=0.0)
        c['precision'] = dict(m['PROPOSAL_PRECISION'])
        c['m_initial'] = int(m['M_INITIAL'])
    raw_u = np.array(data.prior_universal['counts'], float) if c['cp'] else None
    raw_f = np.array(data.prior_inheritance['counts'], float) if (c['cp'] and c['inh']) else None
    st = synthetic.random_state(data, c['Z'], 5, c['inh'], with_source=c['src'], seed=seed + 1)
    st['zones'] = data.planted_zones.copy()          # start at the planted zones (the source is resampled anyway)
    if c['src']:
        st['source'] = synthetic.random_state(data, c['Z'], 5, c['inh'], True, seed=seed + 1, zones=st['zones'])['source']
    return c, data, raw_u, raw_f, st


def run_reference(c, data, st, seed):
    model_cfg = {'N_AREAS': c['Z'], 'MIN_M': 3, 'MAX_M': c.get('max_m', 30), 'INHERITANCE': c['inh'],
                 'SAMPLE_SOURCE': c['src'], 'PRIOR': prior_config(c)}
    model = ref.model.Model(data, model_cfg)
    ops = so.default_operators(c['inh'], c['src'], c['steps'])
    init = ref.zone_sampling.Sample(zones=st['zones'].copy(), weights=st['weights'].copy(),
                                    p_global=st['p_global'].copy(), p_zones=st['p_zones'].copy(),
                                    p_families=None if st['p_families'] is None else st['p_families'].copy())
    random.seed(seed)
    np.random.seed(seed)
    mc3 = {}
    if c.get('mc3'):
        mc3 = dict(mc3=True, swap_period=c['mc3'][0], chain_swaps=c['mc3'][1])
    sampler = ref.zone_sampling.ZoneMCMC(data=data, model=model, n_chains=c.get('n_chains', 1), initial_sample=init,
                                         operators=ops, var_proposal=c['precision'], p_grow_connected=0.85,
                                         initial_size=c['m_initial'], **mc3)
    t0 = time.perf_counter()
    sampler.generate_samples(c['n_steps'], c['n_samples'])
    dt = time.perf_counter() - t0
    s = sampler.statistics
    # quirk Q5: the logged arrays are references that later in-place Gibbs updates may have changed; fine for means
    out = summarise(s['sample_zones'], s['sample_weights'], s['sample_p_global'], s['sample_likelihood'],
                    s['sample_prior'], burn=c['n_samples'] // 5, planted=data.planted_zones.any(0))
    names = sorted(ops)
    out['acc'] = np.array([s['accept_operator'].get(k, 0) / max(1, s['accept_operator'].get(k, 0) +
                                                               s['reject_operator'].get(k, 0)) for k in names])
    out['swap_ratio'] = np.float64(s['swap_ratio'] if c.get('mc3') else 0.0)
    return out, names, c['n_steps'] * c.get('n_chains', 1) / dt, model


def run_oracle(c, data, st, seed, gibbs_g, alpha_g, gibbs_f, alpha_f, n_steps=None):
    ops = so.default_operators(c['inh'], c['src'], c['steps'])
    smp = so.OracleSampler(data.features, data.states, data.families, data.network['adj_mat'], c['inh'], c['src'],
                           prior_config(c), ops, c['precision'], 3, c.get('max_m', 30), 0.85, st, seed=seed,
                           cost_mat=data.geo_prior['cost_matrix'], gibbs_counts_global=gibbs_g, alpha_global=alpha_g,
                           gibbs_counts_fam=gibbs_f, alpha_fam=alpha_f)
    if c['src']:
        smp.gibbs_sample_sources(smp.st)          # generate_initial_sample resamples the sources (:1228-1231)
        smp.st.everything_changed()
        smp.ll, smp.prior = smp.lik(smp.st), smp.pri(smp.st)
    n_steps = n_steps or c['n_steps']
    every = c['n_steps'] // c['n_samples']
    zs, ws, pgs, lls, prs = [], [], [], [], []
    t0 = time.perf_counter()
    for i in range(n_steps):
        smp.step()
        if i % every == 0:
            zs.append(smp.st.zones.copy()); ws.append(smp.st.weights.copy()); pgs.append(smp.st.p_global.copy())
            lls.append(smp.ll); prs.append(smp.prior)
    dt = time.perf_counter() - t0
    out = summarise(zs, ws, pgs, lls, prs, burn=len(zs) // 5, planted=data.planted_zones.any(0))
    names = sorted(ops)
    out['acc'] = np.array([smp.accepted.get(k, 0) / max(1, smp.accepted.get(k, 0) + smp.rejected.get(k, 0))
                           for k in names])
    return out, n_steps / dt


def _ref_job(case, seed):
    import contextlib
    import io
    c, data, raw_u, raw_f, st = build(case)
    with contextlib.redirect_stdout(io.StringIO()):
        out, names, sps, _ = run_reference(c, data, st, seed)
    return out, names, sps


def main(cases):
    os.chdir('/tmp')
    for case in cases:
        c, data, raw_u, raw_f, st = build(case)
        n_runs = c.get('n_runs', N_RUNS)
        with mp.get_context('fork').Pool(min(n_runs, os.cpu_count() or 1)) as pool:
            res = pool.starmap(_ref_job, [(case, 100 + r) for r in range(n_runs)])
        runs, rate, names = [r[0] for r in res], [r[2] for r in res], res[0][1]
        model = ref.model.Model(data, {'N_AREAS': c['Z'], 'MIN_M': 3, 'MAX_M': c.get('max_m', 30), 'INHERITANCE': c['inh'],
                                       'SAMPLE_SOURCE': c['src'], 'PRIOR': prior_config(c)})
        gibbs_g = alpha_g = gibbs_f = alpha_f = None
        if c['cp']:
            gibbs_g, alpha_g = orc.dirichlet_prior_tables(raw_u, c.get('scales', (10., 10.))[0])
            assert np.array_equal(gibbs_g, model.prior.prior_p_global.counts)
            if c['inh']:
                gibbs_f, alpha_f = orc.dirichlet_prior_tables(raw_f, c.get('scales', (10., 10.))[1])
        keys = ['occ_freq', 'zone_size', 'planted_occ', 'w_mean', 'w_comp', 'pg_mean', 'll_mean', 'prior_mean', 'acc',
                'swap_ratio']
        if c.get('skip_port'):
            o_out, o_rate = {k: np.mean([r[k] for r in runs], axis=0) for k in keys}, float('nan')
        else:
            o_out, o_rate = run_oracle(c, data, st, 5, gibbs_g, alpha_g, gibbs_f, alpha_f)
            o_out['swap_ratio'] = np.mean([r['swap_ratio'] for r in runs])
        ref_stack = {k: np.stack([np.asarray(r[k]) for r in runs]) for k in keys}
        print(f'== {case}: reference {np.mean(rate):.0f} steps/s, oracle port {o_rate:.0f} steps/s')
        for k in keys:
            m, s = ref_stack[k].mean(0), ref_stack[k].std(0, ddof=1)
            z = np.abs(np.asarray(o_out[k]) - m) / np.maximum(s * np.sqrt(1 + 1 / n_runs), 1e-3 + 0.02 * np.abs(m))
            print(f'   {k:11s} ref mean {np.ravel(m)[:3]} port {np.ravel(o_out[k])[:3]} max|z| {np.max(z):.2f}')
        adj = data.network['adj_mat'].tocsr()
        save = dict(
            features_idx=orc.state_index(data.features), n_states=np.int64(c['S']), states=data.states,
            families=np.asarray(data.families, bool), adj_indptr=adj.indptr.astype(np.int32),
            adj_indices=adj.indices.astype(np.int32),
            cost=(data.geo_prior['cost_matrix'] if (c['geo'] == 'cost_based' and not c.get('cost_from_locations'))
                  else np.zeros((0, 0))),
            inheritance=np.bool_(c['inh']), sample_source=np.bool_(c['src']), counts_prior=np.bool_(c['cp']),
            geo_type=np.str_(c['geo']), size_type=np.str_(c['size']), geo_scale=np.float64(100.),
            zones=st['zones'], weights=st['weights'], p_global=st['p_global'], p_zones=st['p_zones'],
            n_steps=np.int64(c['n_steps']), n_samples=np.int64(c['n_samples']), op_names=np.array(names),
            max_m=np.int64(c.get('max_m', 30)),
            scale_universal=np.float64(np.nan if c.get('scales', (10., 10.))[0] is None else c.get('scales', (10., 10.))[0]),
            scale_inheritance=np.float64(np.nan if c.get('scales', (10., 10.))[1] is None else c.get('scales', (10., 10.))[1]),
            op_weights=np.array([so.default_operators(c['inh'], c['src'], c['steps'])[k] for k in names]),
            precision=np.array([c['precision'][k] for k in ('weights', 'universal', 'contact', 'inheritance')], float),
            m_initial=np.int64(c['m_initial']), n_chains=np.int64(c.get('n_chains', 1)),
            swap_period=np.int64(c['mc3'][0] if c.get('mc3') else 0),
            chain_swaps=np.int64(c['mc3'][1] if c.get('mc3') else 0),
            planted=data.planted_zones.any(0),
            ref_steps_per_s=np.array(rate), port_steps_per_s=np.float64(o_rate),
            **{'ref_' + k: v for k, v in ref_stack.items()})
        if c.get('cost_from_locations'):
            # the cost matrix is the Euclidean distance matrix of the sites (simulation.py:155-156,
            # preprocessing.py:101-174): the fixture keeps the 951 x 2 locations instead of 951 x 951 doubles
            from scipy.spatial.distance import cdist
            loc = np.asarray(data.network['locations'], float)
            assert np.array_equal(cdist(loc, loc), data.geo_prior['cost_matrix'])
            save['locations'] = loc
        if c['inh']:
            save['p_families'] = st['p_families']
        if c['src']:
            save['source_idx'] = np.argmax(st['source'], axis=-1).astype(np.uint8)
        if c['cp']:
            save['gibbs_counts_global'], save['alpha_global'] = gibbs_g, alpha_g
            if c['inh']:
                save['gibbs_counts_fam'], save['alpha_fam'] = gibbs_f, alpha_f
        np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', case + '.npz'), **save)


if __name__ == '__main__':
    main(sys.argv[1:] or list(CASES))
