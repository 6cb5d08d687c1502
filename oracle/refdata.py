"""TEST INFRASTRUCTURE ONLY -- the reference's BUNDLED experiments (experiments/balkan, experiments/south_america)
loaded with the reference's OWN loaders (util.read_features_from_csv util.py:345-414, preprocessing.compute_network
preprocessing.py:101-174, read_universal_counts / read_inheritance_counts preprocessing.py) into the `data` duck type
of the hot path.  Only usable in the build container (needs /root/reference); the
golden generators store the encoded arrays in tests/golden/*.npz so that nothing here is needed on the GPU box.
"""
import contextlib
import io
import json
import os
import types

import numpy as np

import refshim

EXPERIMENTS = os.path.join(refshim.REFERENCE_ROOT, 'experiments')
# name -> (experiment directory, {prior: scale_counts} as in its config.json)
BUNDLED = {
    'balkan': 'balkan',
    'south_america': 'south_america',
}


def load_bundled(name):
    """-> (data, info).  info: inheritance, scale_universal, scale_inheritance, mcmc (the experiment's mcmc config)."""
    refshim.install()
    import sbayes.preprocessing as rp
    import sbayes.util as ru
    base = os.path.join(EXPERIMENTS, BUNDLED[name])
    cfg = json.load(open(os.path.join(base, 'config.json')))
    dcfg = {k.upper(): v for k, v in cfg['data'].items()}
    ff, fs = os.path.join(base, dcfg['FEATURES']), os.path.join(base, dcfg['FEATURE_STATES'])
    with contextlib.redirect_stdout(io.StringIO()):
        sites, site_names, features, feature_names, state_names, states, families, family_names, _log = \
            ru.read_features_from_csv(ff, fs)
        # CRS = None (Appendix B.2): the adjacency is the Delaunay graph of the projected x / y either way
        network = rp.compute_network(sites)
        prior = cfg['model']['PRIOR']
        pu = prior['universal']
        counts_u, _ = rp.read_universal_counts(feature_names, state_names, os.path.join(base, pu['file']),
                                               pu['file_type'], fs)
        pi = prior['inheritance']
        files = {k: os.path.join(base, v) for k, v in pi['files'].items()}
        counts_i, _ = rp.read_inheritance_counts(family_names, feature_names, state_names, files, pi['file_type'], fs)
    data = types.SimpleNamespace(
        features=np.asarray(features, bool), families=np.asarray(families, bool), states=np.asarray(states, bool),
        network=network, prior_universal={'counts': np.array(counts_u, float)},
        prior_inheritance={'counts': np.array(counts_i, float)},
        geo_prior={'cost_matrix': network['dist_mat']}, state_names=state_names, feature_names=feature_names,
        family_names=family_names, site_names=site_names, is_simulated=False, crs=None,
        planted_zones=np.zeros((0, features.shape[0]), bool))
    info = {'inheritance': bool(cfg['model']['INHERITANCE']), 'scale_universal': pu.get('scale_counts'),
            'scale_inheritance': pi.get('scale_counts'), 'mcmc': cfg['mcmc'], 'model': cfg['model']}
    return data, info


def load_simulated(name='sim_exp1', seed=0, i_contact=3, e_contact=0.5, area=4, inheritance=False):
    """experiments/simulation/<name>: the bundled sites with features drawn by the reference's OWN simulator
    (sbayes/simulation.py:Simulation.run_simulation, seeded), as `run_sim_exp1.py` does for one (strength, area) setup."""
    import pathlib
    import random
    refshim.install()
    from sbayes.simulation import Simulation
    base = os.path.join(EXPERIMENTS, 'simulation', name)
    cfg = json.load(open(os.path.join(base, 'config.json')))
    sim_cfg = json.load(open(os.path.join(refshim.REFERENCE_ROOT, 'sbayes', 'config', 'default_config_simulation.json')))
    sim_cfg.update(cfg['simulation'])
    sim_cfg.update(SITES=os.path.join(base, cfg['simulation']['SITES']))
    if i_contact is not None:       # sim_exp1 leaves these to its run script; sim_exp2 / 3 fix them in config.json
        sim_cfg.update(I_CONTACT=i_contact, E_CONTACT=e_contact, AREA=area)
    exp = types.SimpleNamespace(path_results=pathlib.Path('/tmp'), config={'simulation': sim_cfg})
    random.seed(seed)
    np.random.seed(seed)
    sim = Simulation(experiment=exp)
    with contextlib.redirect_stdout(io.StringIO()):
        sim.run_simulation()
    n = sim.features.shape[0]
    data = types.SimpleNamespace(
        features=np.asarray(sim.features, bool),
        families=np.asarray(sim.families, bool) if inheritance else np.zeros((0, n), bool),
        states=np.asarray(sim.states, bool),
        network=sim.network, prior_universal={}, prior_inheritance={}, geo_prior=sim.geo_prior,
        state_names=sim.state_names, feature_names=sim.feature_names, family_names=None, site_names=sim.site_names,
        is_simulated=True, crs=None, planted_zones=np.asarray(sim.areas, bool))
    info = {'inheritance': bool(inheritance), 'scale_universal': None, 'scale_inheritance': None, 'mcmc': cfg['mcmc'],
            'model': cfg['model']}
    return data, info
