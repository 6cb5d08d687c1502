"""TEST INFRASTRUCTURE ONLY -- golden vectors for the callers and file formats either side of the hot path
(SURVEY.md 8f rows 1-2): `contribution_per_area` (postprocessing.py:271-313), `match_areas` (:205-268), `rank_areas`
(:316-365), `MCMC.eval_ground_truth` (mcmc_setup.py:122-172) and the result writer `samples2file` (util.py:846-907).

Run in the build container (where /root/reference exists):   python oracle/make_golden_results.py
For every case the UNMODIFIED reference's `MCMC` facade (mcmc_setup.py:18-234) runs a short chain, evaluates the
per-area contributions, matches / ranks the areas and writes its result files; the fixture `tests/golden/results_*.npz`
keeps
    the `statistics` dict before matching (logged states, ll / prior, single-area lh / prior / posterior),
    the same lists after match_areas + rank_areas,
    the BYTES of stats_*.txt / areas_*.txt (and ground_truth/stats.txt / areas.txt for simulated data),
    the data arrays needed to rebuild the `data` object on the GPU box.
"""
import contextlib
import copy
import io
import json
import logging
import os
import pathlib
import random
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import refdata  # noqa: E402
import refshim  # noqa: E402
import sbayes_oracle as orc  # noqa: E402

refshim.install()
import sbayes.mcmc_setup as ref_setup  # noqa: E402

MCMC_CFG = {'N_STEPS': 400, 'N_SAMPLES': 8, 'N_RUNS': 1, 'N_CHAINS': 1, 'P_GROW_CONNECTED': 0.85, 'M_INITIAL': 4,
            'PROPOSAL_PRECISION': {'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20},
            'STEPS': {'area': 0.3, 'weights': 0.3, 'universal': 0.05, 'contact': 0.3, 'inheritance': 0.05, 'source': 0.0},
            'WARM_UP': {'N_WARM_UP_STEPS': 50, 'N_WARM_UP_CHAINS': 2}}


def case_balkan():
    data, info = refdata.load_bundled('balkan')
    prior = {'geo': {'type': 'cost_based', 'scale': 200.}, 'area_size': {'type': 'uniform'},
             'weights': {'type': 'uniform'}, 'contact': {'type': 'uniform'},
             'universal': {'type': 'counts', 'scale_counts': info['scale_universal']},
             'inheritance': {'type': 'counts', 'scale_counts': info['scale_inheritance']}}
    cfg = {'model': {'N_AREAS': 3, 'MIN_M': 3, 'MAX_M': 8, 'INHERITANCE': True, 'SAMPLE_SOURCE': True, 'PRIOR': prior},
           'mcmc': copy.deepcopy(MCMC_CFG), 'results': {'FILE_INFO': 'n'}}
    return data, cfg, {'scale_universal': info['scale_universal'], 'scale_inheritance': info['scale_inheritance']}


def case_simexp1():
    import sbayes.simulation as rsim
    base = os.path.join(refdata.EXPERIMENTS, 'simulation', 'sim_exp1')
    scfg = json.load(open(os.path.join(refshim.REFERENCE_ROOT, 'sbayes', 'config', 'default_config_simulation.json')))
    scfg.update(json.load(open(os.path.join(base, 'config.json')))['simulation'])
    scfg.update(SITES=os.path.join(base, 'data', 'sites.csv'), I_CONTACT=3, E_CONTACT=0.5, AREA=4, STRENGTH=1)
    exp = types.SimpleNamespace(path_results=pathlib.Path('/tmp'), config={'simulation': scfg})
    random.seed(11)
    np.random.seed(11)
    sim = rsim.Simulation(experiment=exp)
    with contextlib.redirect_stdout(io.StringIO()):
        sim.run_simulation()
    prior = {'geo': {'type': 'uniform'}, 'area_size': {'type': 'none'}, 'weights': {'type': 'uniform'},
             'contact': {'type': 'uniform'}, 'universal': {'type': 'uniform'}}
    cfg = {'model': {'N_AREAS': 1, 'MIN_M': 3, 'MAX_M': 200, 'INHERITANCE': False, 'SAMPLE_SOURCE': True, 'PRIOR': prior},
           'mcmc': copy.deepcopy(MCMC_CFG), 'results': {'FILE_INFO': 's'},
           'simulation': {'INHERITANCE': bool(scfg['INHERITANCE']), 'STRENGTH': 1, 'AREA': 4}}
    cfg['mcmc']['M_INITIAL'] = 10
    # experiment_setup.py:223-232 zeroes the inheritance steps of a model without inheritance
    cfg['mcmc']['STEPS'] = {'area': 0.3, 'weights': 0.3, 'universal': 0.1, 'contact': 0.3, 'inheritance': 0.0, 'source': 0.0}
    return sim, cfg, {}


CASES = {'results_balkan': case_balkan, 'results_simexp1': case_simexp1}


def stack(lst):
    return np.stack([np.asarray(a) for a in lst])


def main(cases):
    os.chdir('/tmp')
    for name in cases:
        data, cfg, extra = CASES[name]()
        tmp = pathlib.Path(tempfile.mkdtemp(prefix='sb_results_'))
        exp = types.SimpleNamespace(config=cfg, path_results=tmp, logger=logging.getLogger('golden'))
        random.seed(5)
        np.random.seed(5)
        raw_u = np.array(data.prior_universal['counts'], float) if cfg['model']['PRIOR']['universal']['type'] == 'counts' else None
        raw_f = (np.array(data.prior_inheritance['counts'], float)
                 if cfg['model']['PRIOR'].get('inheritance', {}).get('type') == 'counts' else None)
        with contextlib.redirect_stdout(io.StringIO()):
            mc = ref_setup.MCMC(data, exp)
            mc.sample()
            pre = copy.deepcopy(mc.samples)
            mc.save_samples(run=1)
        post = mc.samples
        inh = cfg['model']['INHERITANCE']
        files = {}
        for p in sorted(tmp.rglob('*.txt')):
            files[str(p.relative_to(tmp))] = np.frombuffer(p.read_bytes(), dtype=np.uint8)
        adj = data.network['adj_mat'].tocsr()
        adj.sort_indices()
        save = dict(
            features_idx=orc.state_index(np.asarray(data.features, bool)), n_states=np.int64(data.features.shape[2]),
            states=np.asarray(data.states, bool), families=np.asarray(data.families, bool) if inh else np.zeros((0, data.features.shape[0]), bool),
            adj_indptr=adj.indptr.astype(np.int32), adj_indices=adj.indices.astype(np.int32),
            locations=np.asarray(data.network['locations'], float),
            config=np.str_(json.dumps(cfg)), is_simulated=np.bool_(data.is_simulated),
            names=np.str_(json.dumps({'feature_names': data.feature_names, 'state_names': data.state_names,
                                      'family_names': data.family_names if inh else {'external': [], 'internal': []}},
                                     default=lambda o: o.tolist() if hasattr(o, 'tolist') else str(o))),
            file_names=np.array(list(files)), **{f'file_{i}': v for i, v in enumerate(files.values())})
        if cfg['model']['PRIOR']['geo']['type'] == 'cost_based':
            from scipy.spatial.distance import cdist
            assert np.array_equal(cdist(save['locations'], save['locations']), data.geo_prior['cost_matrix'])
        if raw_u is not None:
            save['raw_counts_universal'] = raw_u
            save['scale_universal'] = np.float64(np.nan if extra['scale_universal'] is None else extra['scale_universal'])
        if raw_f is not None:
            save['raw_counts_inheritance'] = raw_f
            save['scale_inheritance'] = np.float64(np.nan if extra['scale_inheritance'] is None else extra['scale_inheritance'])
        for tag, smp in (('pre', pre), ('post', post)):
            save[f'{tag}_zones'] = stack(smp['sample_zones']).astype(bool)
            save[f'{tag}_p_zones'] = stack(smp['sample_p_zones'])
            for k in ('lh', 'prior', 'posterior'):
                save[f'{tag}_{k}_single'] = np.asarray(smp[f'sample_{k}_single_zones'], float)
        save['weights'] = stack(pre['sample_weights'])
        save['p_global'] = stack(pre['sample_p_global'])
        if inh:
            save['p_families'] = stack(pre['sample_p_families'])
        save['likelihood'] = np.asarray(pre['sample_likelihood'], float)
        save['prior'] = np.asarray(pre['sample_prior'], float)
        if data.is_simulated:
            save.update(true_zones=np.asarray(data.areas, bool), true_weights=np.asarray(post['true_weights'], float),
                        data_weights=np.asarray(data.weights, float),
                        true_p_global=np.asarray(post['true_p_global'], float), true_p_zones=np.asarray(post['true_p_zones'], float),
                        true_ll=np.float64(post['true_ll']), true_prior=np.float64(post['true_prior']),
                        true_lh_single=np.asarray(post['true_lh_single_zones'], float),
                        true_prior_single=np.asarray(post['true_prior_single_zones'], float))
            if data.p_inheritance is not None:
                save['true_p_families'] = np.asarray(data.p_inheritance, float)
        out = os.path.join(ROOT, 'tests', 'golden', name + '.npz')
        np.savez_compressed(out, **save)
        print(f'== {name}: {len(pre["sample_zones"])} samples, files {list(files)}, {os.path.getsize(out)} B')


if __name__ == '__main__':
    main(sys.argv[1:] or list(CASES))
