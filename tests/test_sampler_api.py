"""GPU: `ZoneMCMC` / `ZoneMCMCWarmup` / `MCMC` with the reference's call sequence (mcmc_setup.py:96-118, 174-185):
warm-up -> best sample -> sampling -> statistics -> contribution_per_area -> match / rank -> result files."""
import logging
import pathlib
from types import SimpleNamespace

import numpy as np
import pytest

from helpers import PRECISION, case_data, case_model_config, load_case, load_run

pytestmark = pytest.mark.gpu

STAT_KEYS = ['sample_id', 'sample_likelihood', 'sample_prior', 'sample_zones', 'sample_weights', 'sample_p_global',
             'sample_p_zones', 'sample_p_families', 'last_sample', 'acceptance_ratio', 'accepted_steps', 'n_swaps',
             'accepted_swaps', 'swap_ratio', 'accept_operator', 'reject_operator', 'sampling_time', 'time_per_sample']


def _mcmc_config(n_steps, n_samples, n_chains=1, mc3=None):
    c = {'N_STEPS': n_steps, 'N_SAMPLES': n_samples, 'N_RUNS': 1, 'P_GROW_CONNECTED': 0.85,
         'PROPOSAL_PRECISION': dict(PRECISION), 'M_INITIAL': 5, 'N_CHAINS': n_chains, 'SEED': 123,
         'STEPS': {'area': 0.05, 'weights': 0.4, 'universal': 0.05, 'contact': 0.4, 'inheritance': 0.1, 'source': 0.0},
         'WARM_UP': {'N_WARM_UP_STEPS': 300, 'N_WARM_UP_CHAINS': 6}}
    if mc3:
        c['MC3'] = mc3
    return c


@pytest.mark.parametrize('sample_source', [True, False])
def test_zone_mcmc_generate_samples(sample_source, built_lib):
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import Sample, ZoneMCMC, ZoneMCMCWarmup
    c = load_case('C_samerica_like')
    data = case_data(c)
    model = Model(data, case_model_config(c, sample_source=sample_source))
    kind = 'gibbs_sample' if sample_source else 'alter'
    ops = {'shrink_zone': 0.02, 'grow_zone': 0.02, 'swap_zone': 0.01, 'gibbsish_sample_zones': 0.0,
           f'{kind}_weights': 0.4, f'{kind}_p_global': 0.05, f'{kind}_p_zones': 0.4, f'{kind}_p_families': 0.1}
    kw = dict(data=data, model=model, operators=ops, var_proposal=PRECISION, p_grow_connected=0.85, initial_size=5,
              seed=7)
    warm = ZoneMCMCWarmup(n_chains=6, **kw)
    assert len(warm.max_size) == 6 and warm.max_size[0] <= warm.max_size[-1] <= model.max_size
    assert set(warm.p_grow_connected) <= {0.95, 0.85}
    best = warm.generate_samples(n_steps=0, n_samples=0, warm_up=True, warm_up_steps=300)
    warm.close()
    assert isinstance(best, Sample) and best.zones.shape == (c['Z'], c['N'])
    assert (best.zones.sum(axis=0) <= 1).all() and (best.zones.sum(axis=1) >= model.min_size).all()
    assert (best.source is not None) == sample_source

    n_steps, n_samples = 2000, 20
    smp = ZoneMCMC(n_chains=4, initial_sample=best, mc3=True, swap_period=100, chain_swaps=2, **kw)
    assert smp.generate_samples(n_steps, n_samples) is None
    st = smp.statistics
    for k in STAT_KEYS:
        assert k in st, k
    assert st['sample_id'] == list(range(n_samples))
    assert len(st['sample_zones']) == n_samples and st['sample_zones'][0].shape == (c['Z'], c['N'])
    assert st['sample_weights'][0].shape == (c['F'], c['C']) and st['sample_p_global'][0].shape == (1, c['F'], c['S'])
    assert st['n_swaps'] == (n_steps // 100) * 2 and 0 <= st['accepted_swaps'] <= st['n_swaps']
    total = sum(st['accept_operator'].values()) + sum(st['reject_operator'].values())
    assert total == n_steps * 4                                       # every chain, every step (quirk Q4)
    assert st['accepted_steps'] == sum(st['accept_operator'].values())
    assert isinstance(st['last_sample'], Sample)
    # the logged ll / prior are the model's values of the logged states (1e-9)
    for i in (0, n_samples - 1):
        s = Sample(st['sample_zones'][i], st['sample_weights'][i], st['sample_p_global'][i], st['sample_p_zones'][i],
                   st['sample_p_families'][i])
        assert abs(model.prior(s) - st['sample_prior'][i]) <= 1e-9 * max(1., abs(st['sample_prior'][i]))
        if not sample_source:
            assert abs(model.likelihood(s) - st['sample_likelihood'][i]) <= 1e-9 * abs(st['sample_likelihood'][i])
    last = st['last_sample']
    c0 = smp.chain_idx[0]
    assert abs(model.likelihood(last) - smp._ll[c0]) <= 1e-9 * abs(smp._ll[c0])
    smp.close()


def test_mcmc_facade_writes_reference_files(built_lib, tmp_path):
    from contact_zones_b200.mcmc_setup import MCMC
    from contact_zones_b200.results import parse_area_columns
    c = load_case('B_balkan_like')
    data = case_data(c)
    cfg = {'model': case_model_config(c), 'mcmc': _mcmc_config(1000, 10), 'results': {'FILE_INFO': 'n'}}
    exp = SimpleNamespace(config=cfg, path_results=pathlib.Path(tmp_path), logger=logging.getLogger('t'))
    mc = MCMC(data, exp)
    assert abs(sum(mc.ops.values()) - 1.0) < 1e-12 and mc.ops['gibbsish_sample_zones'] == 0.0
    mc.warm_up()
    mc.sample()
    assert len(mc.samples['sample_lh_single_zones']) == 10 and len(mc.samples['sample_lh_single_zones'][0]) == c['Z']
    mc.save_samples(run=1)
    mc.log_statistics()
    fi = f"n{c['Z']}"
    stats = (tmp_path / fi / f'stats_{fi}_1.txt').read_text().splitlines()
    areas = (tmp_path / fi / f'areas_{fi}_1.txt').read_text().splitlines()
    hdr = stats[0].split('\t')
    assert hdr[:4] == ['Sample', 'posterior', 'likelihood', 'prior'] and hdr[4] == 'size_a0'
    assert 'w_universal_f0' in hdr and 'alpha_f0_s0' in hdr and 'gamma_a1_f0_s0' in hdr and 'beta_fam0_f0_s0' in hdr
    assert hdr[-3:] == [f"lh_a{c['Z']}", f"prior_a{c['Z']}", f"post_a{c['Z']}"]
    assert len(stats) == 11 and len(areas) == 10
    row = dict(zip(hdr, stats[1].split('\t')))
    assert abs(float(row['posterior']) - float(row['likelihood']) - float(row['prior'])) < 1e-6
    z = parse_area_columns(areas[0])
    assert z.shape == (c['Z'], c['N']) and int(row['size_a0']) == z[0].sum()


def test_seed_is_expanded_per_run_and_phase(built_lib, tmp_path):
    """Same SEED: the same run index reproduces the files byte for byte, another run index gives another chain, and
    the sampling phase does not replay the warm-up's streams (sub-seeds per sampler instance)."""
    from contact_zones_b200.mcmc_setup import MCMC
    c = load_case('B_balkan_like')
    data = case_data(c)
    outs = []
    for tag, run in (('a', 0), ('b', 0), ('c', 1)):
        cfg = {'model': case_model_config(c), 'mcmc': _mcmc_config(600, 6), 'results': {'FILE_INFO': 'n'}}
        d = tmp_path / tag
        d.mkdir()
        mc = MCMC(data, SimpleNamespace(config=cfg, path_results=pathlib.Path(d), logger=logging.getLogger('t')), run=run)
        mc.warm_up()
        mc.sample()
        mc.save_samples(run=run)
        mc.sampler.close()
        fi = f"n{c['Z']}"
        outs.append((d / fi / f'stats_{fi}_{run}.txt').read_bytes())
        assert mc._sampler_kwargs(0)['seed'] != mc._sampler_kwargs(1)['seed']
    assert outs[0] == outs[1] and outs[0] != outs[2]


def test_cli_end_to_end(built_lib, tmp_path):
    """`python -m contact_zones_b200 config.json` on a CSV dataset: counts priors, cost-based geo prior from a cost
    file, inheritance, warm-up, sampling, result files (sbayes/cli.py:30-84)."""
    from test_data_loading import write_dataset
    from contact_zones_b200.cli import main
    from contact_zones_b200.results import parse_area_columns
    write_dataset(tmp_path, n=60, seed=3)
    last = main(config=tmp_path / 'config.json', experiment_name='cli_run', custom_settings={'mcmc': {'SEED': 4}})
    out = tmp_path / 'results' / 'cli_run'
    assert (out / 'experiment.log').exists()
    stats = (out / 'n2' / 'stats_n2_0.txt').read_text().splitlines()
    areas = (out / 'n2' / 'areas_n2_0.txt').read_text().splitlines()
    assert len(stats) == 5 and len(areas) == 4
    hdr = stats[0].split('\t')
    assert 'w_inheritance_f_bin' in hdr and 'alpha_f_tri_c' in hdr and 'gamma_a2_f_four_z' in hdr and 'beta_Beta_f_col_red' in hdr
    z = parse_area_columns(areas[-1])
    assert z.shape == (2, 60) and (z.sum(axis=1) >= 3).all() and (z.sum(axis=0) <= 1).all()
    vals = [dict(zip(hdr, r.split('\t'))) for r in stats[1:]]
    assert all(np.isfinite(float(v['likelihood'])) and np.isfinite(float(v['prior'])) for v in vals)
    assert float(vals[0]['prior']) != 0.0                                  # counts + cost-based priors are active
    assert last.zones.shape == (2, 60) and last.source.shape == (60, 4, 3)


@pytest.mark.parametrize('seed', range(12))
def test_api_on_random_shapes(seed, built_lib):
    """The host mirror (Model, ZoneMCMCWarmup, ZoneMCMC) on seeded random synthetic data sets, the reference's call
    sequence: warm-up -> best sample -> sampling; the logged values are the model's values of the logged states."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import Sample, ZoneMCMC, ZoneMCMCWarmup
    rng = np.random.default_rng(500 + seed)
    z = int(rng.integers(1, 4))
    n = int(rng.integers(25, 200))
    f, s_ = int(rng.integers(2, 40)), int(rng.integers(2, 7))
    fam = int(rng.integers(0, 6))
    inh = fam > 0
    sample_source = bool(rng.integers(0, 2))
    geo = bool(rng.integers(0, 2))
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=seed, with_dist=True, ragged_states=bool(rng.integers(0, 2)),
                               na_frac=0.05, counts_prior=bool(rng.integers(0, 2)))
    has_counts = bool(data.prior_universal)
    prior = {'geo': {'type': 'cost_based', 'scale': 150.} if geo else {'type': 'uniform'},
             'area_size': {'type': ['none', 'uniform', 'quadratic'][int(rng.integers(0, 3))]},
             'weights': {'type': 'uniform'}, 'contact': {'type': 'uniform'},
             'universal': {'type': 'counts', 'scale_counts': 10., 'file': 'x'} if has_counts else {'type': 'uniform'}}
    if inh:
        prior['inheritance'] = ({'type': 'counts', 'scale_counts': 10., 'files': {}} if has_counts
                                else {'type': 'uniform'})
    max_m = int(rng.integers(8, 25))
    cfg = {'N_AREAS': z, 'MIN_M': 3, 'MAX_M': max_m, 'INHERITANCE': inh, 'SAMPLE_SOURCE': sample_source, 'PRIOR': prior}
    model = Model(data, cfg)
    kind = 'gibbs_sample' if sample_source else 'alter'
    ops = {'shrink_zone': 0.04, 'grow_zone': 0.04, 'swap_zone': 0.02, 'gibbsish_sample_zones': 0.0,
           f'{kind}_weights': 0.4, f'{kind}_p_global': 0.1, f'{kind}_p_zones': 0.3}
    if inh:
        ops[f'{kind}_p_families'] = 0.1
    kw = dict(data=data, model=model, operators=ops, var_proposal=PRECISION, p_grow_connected=0.85, initial_size=4,
              seed=seed)
    warm = ZoneMCMCWarmup(n_chains=int(rng.integers(1, 6)), **kw)
    best = warm.generate_samples(n_steps=0, n_samples=0, warm_up=True, warm_up_steps=200)
    warm.close()
    assert (best.zones.sum(axis=0) <= 1).all() and (best.zones.sum(axis=1) >= 3).all()
    n_chains = int(rng.integers(1, 4))
    smp = ZoneMCMC(n_chains=n_chains, initial_sample=best, mc3=n_chains > 1, swap_period=50, chain_swaps=1, **kw)
    smp.generate_samples(600, 6)
    st = smp.statistics
    assert len(st['sample_zones']) == 6 and st['sample_zones'][0].shape == (z, n)
    for i in (0, 5):
        s = Sample(st['sample_zones'][i], st['sample_weights'][i], st['sample_p_global'][i], st['sample_p_zones'][i],
                   st['sample_p_families'][i])
        assert abs(model.prior(s) - st['sample_prior'][i]) <= 1e-9 * max(1., abs(st['sample_prior'][i]))
        if not sample_source:
            assert abs(model.likelihood(s) - st['sample_likelihood'][i]) <= 1e-9 * abs(st['sample_likelihood'][i])
    last = st['last_sample']
    c0 = smp.chain_idx[0]
    assert abs(model.likelihood(last) - smp._ll[c0]) <= 1e-9 * abs(smp._ll[c0])
    smp.close()


@pytest.mark.parametrize('seed', range(6))
def test_mcmc_facade_on_random_shapes(seed, built_lib, tmp_path):
    """`MCMC` end to end (warm-up, sampling, contribution_per_area, match / rank, result files) on seeded random data
    sets: Z = 1..4 zones, with / without families, MC3 on / off; the files parse back to consistent values."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.mcmc_setup import MCMC
    from contact_zones_b200.results import parse_area_columns
    rng = np.random.default_rng(900 + seed)
    z, n = int(rng.integers(1, 5)), int(rng.integers(30, 150))
    f, s_, fam = int(rng.integers(2, 30)), int(rng.integers(2, 6)), int(rng.integers(0, 5))
    inh = fam > 0
    data = synthetic.make_data(n, f, s_, n_families=fam, seed=seed, with_dist=True, ragged_states=True, na_frac=0.05)
    prior = {'geo': {'type': 'cost_based', 'scale': 100.} if seed % 2 else {'type': 'uniform'},
             'area_size': {'type': 'uniform'}, 'weights': {'type': 'uniform'}, 'contact': {'type': 'uniform'},
             'universal': {'type': 'uniform'}}
    if inh:
        prior['inheritance'] = {'type': 'uniform'}
    model_cfg = {'N_AREAS': z, 'MIN_M': 3, 'MAX_M': 15, 'INHERITANCE': inh, 'SAMPLE_SOURCE': bool(seed % 3), 'PRIOR': prior}
    n_chains = 1 + seed % 3
    mcmc_cfg = _mcmc_config(600, 6, n_chains=n_chains,
                            mc3={'ACTIVATE': True, 'N_CHAINS': n_chains, 'SWAP_PERIOD': 100, 'N_SWAPS': 1} if n_chains > 1 else None)
    if not inh:
        # (the Experiment sets the inheritance steps to 0 when inheritance is disabled, experiment_setup.py:223-227)
        mcmc_cfg['STEPS'] = {'area': 0.1, 'weights': 0.4, 'universal': 0.1, 'contact': 0.4, 'inheritance': 0.0, 'source': 0.0}
    cfg = {'model': model_cfg, 'mcmc': mcmc_cfg, 'results': {'FILE_INFO': 'n'}}
    exp = SimpleNamespace(config=cfg, path_results=pathlib.Path(tmp_path), logger=logging.getLogger('t'))
    mc = MCMC(data, exp)
    mc.warm_up()
    mc.sample()
    mc.save_samples(run=0)
    mc.log_statistics()
    fi = f'n{z}'
    stats = (tmp_path / fi / f'stats_{fi}_0.txt').read_text().splitlines()
    areas = (tmp_path / fi / f'areas_{fi}_0.txt').read_text().splitlines()
    assert len(stats) == 7 and len(areas) == 6
    hdr = stats[0].split('\t')
    for line, arow in zip(stats[1:], areas):
        row = dict(zip(hdr, line.split('\t')))
        assert abs(float(row['posterior']) - float(row['likelihood']) - float(row['prior'])) < 1e-6
        zz = parse_area_columns(arow)
        assert zz.shape == (z, n) and (zz.sum(axis=0) <= 1).all() and (zz.sum(axis=1) >= 3).all()
        assert all(int(row[f'size_a{k}']) == zz[k].sum() for k in range(z))
