"""GPU: distributional parity of the CUDA sampler (persistent kernel K8 + operators K7, through the C ABI) with the
reference's CPU sampler.  BASELINE.json north_star: "because the RNG differs, posterior summaries ... must agree with
the reference's CPU run within Monte-Carlo error".  The fixtures hold six independent runs of the UNMODIFIED reference
(oracle/make_golden_runs.py); we run 16 chains of the same length from the same initial state and compare the
label-switching-invariant summaries.  Bar: |z| < 4.5 on every statistic (z against the run-to-run spread of both
sides)."""
import numpy as np
import pytest

from helpers import PRECISION, load_run, run_cases, run_z_scores, summarise_run, vector_z_scores

pytestmark = pytest.mark.gpu
N_CHAINS = 16
MC3_CASES = ('run_G_balkan_mc3', 'run_H_samerica_mc3', 'run_I_simexp2')     # run through the ZoneMCMC API below
Z_MAX = 4.5


@pytest.mark.parametrize('name', [n for n in run_cases() if n not in MC3_CASES])
def test_posterior_summaries_match_reference(name, built_lib):
    from contact_zones_b200 import _lib
    from contact_zones_b200.engine import Engine
    c = load_run(name)
    g = c['g']
    eng = Engine(g['features_idx'], g['states'], c['families'] if c['inheritance'] else None, c['adj'], c['cost'],
                 n_zones=c['Z'], n_chains=N_CHAINS, inheritance=c['inheritance'], sample_source=c['sample_source'])
    eng.set_priors(size_type=c['size_type'], geo_type=c['geo_type'], geo_scale=c['geo_scale'],
                   p_global_type='counts' if c['counts_prior'] else 'uniform',
                   p_families_type='counts' if (c['counts_prior'] and c['inheritance']) else 'uniform',
                   gibbs_counts_global=g.get('gibbs_counts_global'), alpha_global=g.get('alpha_global'),
                   gibbs_counts_fam=g.get('gibbs_counts_fam'), alpha_fam=g.get('alpha_fam'))
    for ch in range(N_CHAINS):
        eng.set_state(ch, g['zones'], g['weights'], g['p_global'], g['p_zones'], c['p_families'],
                      g['source_idx'] if c['sample_source'] else None)
    eng.eval()
    eng.set_operators(c['operators'], PRECISION, min_size=3, max_size=int(g.get('max_m', 30)), p_grow_connected=0.85, seed=2024)
    n_steps, n_samples = int(g['n_steps']), int(g['n_samples'])
    every = n_steps // n_samples
    zs = [[] for _ in range(N_CHAINS)]
    ws = [[] for _ in range(N_CHAINS)]
    pgs = [[] for _ in range(N_CHAINS)]
    lls, prs = [], []
    for _ in range(n_samples):
        eng.run(every)
        ll, pr = eng.chain_scalars()
        lls.append(ll)
        prs.append(pr)
        for ch in range(N_CHAINS):
            s = eng.get_state(ch, fields=('zones', 'weights', 'p_global'))
            zs[ch].append(s['zones'])
            ws[ch].append(s['weights'])
            pgs[ch].append(s['p_global'])
    acc, rej = eng.op_stats()
    names = [str(k) for k in g['op_names']]
    lls, prs = np.array(lls), np.array(prs)
    chains = []
    for ch in range(N_CHAINS):
        a = [acc[ch, _lib.OP_IDS[k]] / max(1, acc[ch, _lib.OP_IDS[k]] + rej[ch, _lib.OP_IDS[k]]) for k in names]
        chains.append(summarise_run(zs[ch], ws[ch], lls[:, ch], prs[:, ch], a, g['planted'], burn=n_samples // 5,
                                    p_global=pgs[ch]))
    # the running ll / prior of the sampler equal a from-scratch evaluation of the final states
    ll_new, pr_new, _ = eng.eval()
    assert np.allclose(lls[-1], ll_new, rtol=1e-9) and np.allclose(prs[-1], pr_new, rtol=1e-9, atol=1e-9)
    eng.close()
    z = run_z_scores(c, chains)
    worst = {k: float(np.max(v)) for k, v in z.items()}
    assert max(worst.values()) < Z_MAX, (name, worst, {k: np.mean([s[k] for s in chains], axis=0) for k in z},
                                         {k: g['ref_' + k].mean(0) for k in z})
    # per-site zone membership frequency, per-(feature, component) weight means, per-(feature, state) p_global means
    vz = vector_z_scores(c, chains)
    assert set(vz) == {'occ_freq', 'w_mean', 'pg_mean'}
    for k, (zmax, bar, z2) in vz.items():
        assert zmax < bar and z2 < 3.0, (name, k, zmax, bar, z2)


@pytest.mark.parametrize('name', [n for n in run_cases() if n in MC3_CASES])
def test_full_form_experiments_through_zone_mcmc(name, built_lib):
    """The bundled experiments in the form BASELINE.json configs[1] / configs[2] name them -- Balkan with 3 areas and
    4 MC3 chains, South America with 5 areas and 8 MC3 chains, each with the operator mix / proposal precision /
    MAX_M of its own config.json -- and a sim_exp2-like SAMPLE_SOURCE = false run with the cost-based geo prior on the
    951 bundled sites, through the drop-in API: `Model` + `ZoneMCMC(mc3=True, swap_period, chain_swaps)`
    .generate_samples -> `statistics`, against 8 runs of the unmodified reference with the same call."""
    from helpers import case_data, case_model_config
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import Sample, ZoneMCMC
    c = load_run(name)
    g = c['g']
    data = case_data(c)
    model = Model(data, case_model_config(c, sample_source=c['sample_source'], max_m=int(g['max_m'])))
    n_chains, mc3 = int(g['n_chains']), int(g['swap_period']) > 0
    init = Sample(g['zones'].astype(bool), g['weights'], g['p_global'], g['p_zones'], c['p_families'])
    n_steps, n_samples = int(g['n_steps']), int(g['n_samples'])
    names = [str(k) for k in g['op_names']]
    ops = dict(c['operators'], gibbsish_sample_zones=0.0)
    runs, swap_ratios = [], []
    n_runs = 12
    for r in range(n_runs):
        smp = ZoneMCMC(data=data, model=model, n_chains=n_chains, initial_sample=init, operators=ops,
                       var_proposal=c['precision'], p_grow_connected=0.85, initial_size=int(g['m_initial']),
                       mc3=mc3, swap_period=int(g['swap_period']) or None, chain_swaps=int(g['chain_swaps']) or None,
                       seed=4000 + r)
        smp.generate_samples(n_steps, n_samples)
        st = smp.statistics
        assert len(st['sample_zones']) == n_samples
        acc = [st['accept_operator'][k] / max(1, st['accept_operator'][k] + st['reject_operator'][k]) for k in names]
        runs.append(summarise_run(st['sample_zones'], st['sample_weights'], st['sample_likelihood'], st['sample_prior'],
                                  acc, g['planted'], burn=n_samples // 5, p_global=st['sample_p_global']))
        swap_ratios.append(st['swap_ratio'])
        smp.close()
    z = run_z_scores(c, runs)
    worst = {k: float(np.max(v)) for k, v in z.items()}
    assert max(worst.values()) < Z_MAX, (name, worst, {k: np.mean([s[k] for s in runs], axis=0) for k in z},
                                         {k: g['ref_' + k].mean(0) for k in z})
    for k, (zmax, bar, z2) in vector_z_scores(c, runs).items():
        assert zmax < bar and z2 < 3.0, (name, k, zmax, bar, z2)
    if mc3:
        ref_sw = np.asarray(g['ref_swap_ratio'], float)
        sw = np.asarray(swap_ratios, float)
        zsw = abs(sw.mean() - ref_sw.mean()) / np.sqrt(max(ref_sw.var(ddof=1), 1e-4) / len(ref_sw) +
                                                       max(sw.var(ddof=1), 1e-4) / len(sw))
        assert zsw < Z_MAX, (name, sw.mean(), ref_sw.mean(), zsw)
