"""Mirror of `sbayes/experiment_setup.py`: config loading with the reference's defaults and checks
(`config/default_config.json`, `experiment_setup.py:66-254`), results folder, logger.  Simulation configs are out of
scope (the dataset simulator is not on the path) and are rejected."""
import copy
import json
import logging
import os
import time
import warnings
from pathlib import Path

REQUIRED = '<REQUIRED>'

DEFAULT_CONFIG = {
    'mcmc': {
        'N_STEPS': 1000000, 'N_SAMPLES': 1000, 'N_RUNS': 1, 'P_GROW_CONNECTED': 0.85,
        'PROPOSAL_PRECISION': {'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20},
        'STEPS': {'area': 0.05, 'weights': 0.4, 'universal': 0.05, 'contact': 0.4, 'inheritance': 0.1, 'source': 0.0},
        'M_INITIAL': 5,
        'WARM_UP': {'N_WARM_UP_STEPS': 100000, 'N_WARM_UP_CHAINS': 15},
    },
    'model': {
        'N_AREAS': REQUIRED, 'MIN_M': 3, 'MAX_M': 50, 'INHERITANCE': REQUIRED, 'SAMPLE_SOURCE': True,
        'PRIOR': {'geo': {'type': 'uniform'}, 'area_size': {'type': 'none'}, 'weights': {'type': 'uniform'},
                  'universal': {'type': 'uniform'}, 'inheritance': {'type': 'uniform'}, 'contact': {'type': 'uniform'}},
    },
    'data': {'FEATURES': REQUIRED, 'FEATURE_STATES': REQUIRED},
}


def set_experiment_name():
    """experiment_setup.py: results folder named after the start time."""
    return time.strftime('%Y-%m-%d_%H-%M')


def set_defaults(cfg, default_cfg):
    """Fill every field missing in `cfg` from `default_cfg`, recursively (experiment_setup.py:295-314).

    >>> set_defaults(cfg={0:0, 1:{1:0}, 2:{2:1}}, default_cfg={1:{1:1}, 2:{1:1, 2:2}})
    {0: 0, 1: {1: 0}, 2: {2: 1, 1: 1}}
    """
    for key, dv in default_cfg.items():
        if key not in cfg:
            cfg[key] = copy.deepcopy(dv)
        elif isinstance(dv, dict) and isinstance(cfg[key], dict):
            set_defaults(cfg[key], dv)
    return cfg


def update_recursive(cfg, new_cfg):
    """Overwrite the fields of `cfg` given in `new_cfg`, recursively (experiment_setup.py:317-334).

    >>> update_recursive(cfg={0:0, 1:{1:0}, 2:{2:1}}, new_cfg={1:{1:1}, 2:{1:1, 2:2}})
    {0: 0, 1: {1: 1}, 2: {2: 2, 1: 1}}
    """
    for key, nv in new_cfg.items():
        if key in cfg and isinstance(nv, dict) and isinstance(cfg[key], dict):
            update_recursive(cfg[key], nv)
        else:
            cfg[key] = nv
    return cfg


def iter_items_recursive(cfg):
    """All key-value pairs of the bottom-level dictionaries (experiment_setup.py:337-355)."""
    for key, value in cfg.items():
        if isinstance(value, dict):
            yield from iter_items_recursive(value)
        else:
            yield key, value


class Experiment:
    """`sbayes.experiment_setup.Experiment`: `.config`, `.path_results`, `.logger`, `.experiment_name`."""

    def __init__(self, experiment_name=None, config_file=None, log=True, custom_settings=None):
        self.experiment_name = set_experiment_name() if experiment_name is None else experiment_name
        self.config_file = None
        self.config = {}
        self.base_directory = None
        self.path_results = None
        self.logger = self.init_logger()
        if config_file is not None:
            self.load_config(config_file, custom_settings=custom_settings)
        if log:
            self.log_experiment()

    def load_config(self, config_file, custom_settings=None):
        self.base_directory, self.config_file = self.decompose_config_path(config_file)
        with open(self.config_file, 'r') as f:
            self.config = json.load(f)
        if 'simulation' in self.config:
            raise NotImplementedError('simulation configs need the dataset simulator, which is outside this build')
        set_defaults(self.config, DEFAULT_CONFIG)
        if custom_settings is not None:
            update_recursive(self.config, custom_settings)
        self.verify_config()
        self.path_results = self.fix_relative_path('{path}/{experiment}/'.format(
            path=self.config['results']['RESULTS_PATH'], experiment=self.experiment_name))
        os.makedirs(self.path_results, exist_ok=True)
        self.add_logger_file(self.path_results)

    @staticmethod
    def decompose_config_path(config_path):
        p = Path(config_path).absolute()
        return p.parent, p

    def fix_relative_path(self, path):
        path = Path(path)
        return path if path.is_absolute() else self.base_directory / path

    def is_simulation(self):
        return 'simulation' in self.config

    def verify_priors(self, priors_cfg, inheritance):
        """experiment_setup.py:127-176: required prior sections, counts-file paths relative to the config file."""
        required = ['geo', 'area_size', 'weights', 'universal', 'contact'] + (['inheritance'] if inheritance else [])
        if not inheritance and priors_cfg.get('inheritance') is not None:
            if priors_cfg['inheritance'].get('type', 'uniform') != 'uniform':
                warnings.warn("Inheritance is not considered in the model. PRIOR for inheritance will not be used.")
            priors_cfg['inheritance'] = None
        for key in required:
            if key not in priors_cfg:
                raise NameError(f"Prior '{key}' is not defined in {self.config_file}.")
            prior = priors_cfg[key]
            if 'type' not in prior:
                raise NameError(f"'type' for prior '{key}' is not defined in {self.config_file}.")
            if key == 'geo' and prior['type'] == 'cost_based':
                prior.setdefault('scale', None)
                if prior['scale'] is None:
                    raise NameError("'scale' is required for the cost_based geo prior.")
                if 'file' in prior:
                    prior['file'] = self.fix_relative_path(prior['file'])
            if key in ('universal', 'inheritance') and prior['type'] == 'counts':
                prior.setdefault('scale_counts', None)
                prior.setdefault('file_type', 'counts_file')
                if key == 'universal':
                    if 'file' not in prior:
                        raise NameError(f"'file' for the universal counts prior is not defined in {self.config_file}.")
                    prior['file'] = self.fix_relative_path(prior['file'])
                else:
                    if 'files' not in prior:
                        raise NameError(f"'files' for the inheritance counts prior are not defined in {self.config_file}.")
                    prior['files'] = {fam: self.fix_relative_path(p) for fam, p in prior['files'].items()}

    def verify_config(self):
        cfg = self.config
        for key, value in iter_items_recursive(cfg):
            if value == REQUIRED:
                raise NameError(f'´{key}´ is not defined in {self.config_file}')
        if 'model' not in cfg:
            raise NameError("Information about the model was not found in " + str(self.config_file))
        self.verify_priors(cfg['model']['PRIOR'], inheritance=cfg['model']['INHERITANCE'])
        cfg['model'].setdefault('NEIGHBOR_DIST', 'euclidean')
        if 'mcmc' not in cfg:
            raise NameError("Information about the MCMC setup was not found in " + str(self.config_file))
        if 'MC3' not in cfg['mcmc']:
            cfg['mcmc']['N_CHAINS'] = 1
        else:
            cfg['mcmc'].setdefault('N_CHAINS', cfg['mcmc']['MC3'].get('N_CHAINS', 1))
        if cfg['mcmc']['N_STEPS'] % cfg['mcmc']['N_SAMPLES'] != 0:
            raise ValueError("Non-consistent spacing between samples. Set N_STEPS to be a multiple of N_SAMPLES. ")
        steps = cfg['mcmc']['STEPS']
        if not cfg['model']['INHERITANCE']:
            if steps.get('inheritance', 0) != 0:
                self.logger.warning('STEPS for inheritance was set to 0, because ´inheritance´ is disabled.')
            steps['inheritance'] = 0.0
        if not cfg['model']['SAMPLE_SOURCE']:
            if steps.get('source', 0) != 0:
                self.logger.warning('STEPS for source was set to 0, because ´SAMPLE_SOURCE´ is disabled.')
            steps['source'] = 0.0
        total = sum(steps.values())
        for op in steps:
            steps[op] = steps[op] / total
        cfg.setdefault('results', {})
        cfg['results'].setdefault('RESULTS_PATH', 'results')
        cfg['results'].setdefault('FILE_INFO', 'n')
        if 'data' not in cfg:
            raise NameError("Provide file paths to data.")
        if not isinstance(cfg['data'], dict) or 'cldf_dataset' in cfg['data']:
            raise NotImplementedError('CLDF datasets need pycldf, which is outside this build; give FEATURES and '
                                      'FEATURE_STATES csv files')
        cfg['data'].setdefault('simulated', False)
        for key in ('FEATURES', 'FEATURE_STATES'):
            if not cfg['data'].get(key):
                raise NameError(f"{key} is empty. Provide the file path.")
            cfg['data'][key] = self.fix_relative_path(cfg['data'][key])

    def init_logger(self):
        logger = logging.Logger('sbayesLogger', level=logging.DEBUG)
        logger.addHandler(logging.StreamHandler())
        return logger

    def add_logger_file(self, path_results):
        self.logger.addHandler(logging.FileHandler(filename=path_results / 'experiment.log'))

    def log_experiment(self):
        self.logger.info("Experiment: %s", self.experiment_name)
        self.logger.info("File location for results: %s", self.path_results)
