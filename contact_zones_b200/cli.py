"""`sbayes <config.json>` on the device sampler (mirror of `sbayes/cli.py:11-84`): experiment -> data -> for every
run (and every N_AREAS if a list is given): warm-up, sampling, statistics, result files."""
import argparse
from pathlib import Path

from .data import Data
from .experiment_setup import Experiment
from .mcmc_setup import MCMC


def run_experiment(experiment, data, run):
    mcmc = MCMC(data=data, experiment=experiment, run=run)
    mcmc.log_setup()
    mcmc.warm_up()
    mcmc.sample()
    mcmc.log_statistics()
    mcmc.save_samples(run=run)
    mcmc.sampler.close()
    return mcmc.samples['last_sample']


def main(config=None, experiment_name=None, custom_settings=None):
    if config is None:
        parser = argparse.ArgumentParser(description="An MCMC algorithm to identify contact zones (B200 build)")
        parser.add_argument("config", type=Path, help="The JSON configuration file")
        parser.add_argument("--name", default=None, help="experiment name (= results sub-folder)")
        args = parser.parse_args()
        config, experiment_name = args.config, args.name
    experiment = Experiment(experiment_name=experiment_name, config_file=config, log=True,
                            custom_settings=custom_settings)
    data = Data(experiment=experiment)
    data.load_features()
    data.load_universal_counts()
    data.load_inheritance_counts()
    data.load_geo_cost_matrix()          # the reference CLI omits this call, so its cost_based geo prior cannot run
    data.log_loading()
    last = None
    for run in range(experiment.config['mcmc']['N_RUNS']):
        n_areas = experiment.config['model']['N_AREAS']
        if isinstance(n_areas, list):
            for n in n_areas:
                experiment.config['model']['N_AREAS'] = n
                last = run_experiment(experiment, data, run)
            experiment.config['model']['N_AREAS'] = n_areas
        else:
            assert isinstance(n_areas, int)
            last = run_experiment(experiment, data, run)
    return last


if __name__ == '__main__':
    main()
