"""Drop-in mirror of `sbayes/mcmc_setup.py`: the `MCMC` facade that wires config -> model -> operator table ->
warm-up -> sampling -> per-area contributions -> result files, on the device sampler."""
import logging

import numpy as np

from .model import Model
from .results import contribution_per_area, match_areas, rank_areas, samples2file
from .sampling import Sample, ZoneMCMC, ZoneMCMCWarmup


class MCMC:
    """mcmc_setup.py:18-234.  `experiment` needs `.config`, `.path_results` (pathlib.Path) and `.logger`."""

    def __init__(self, data, experiment, run=0):
        self.data = data
        self.run = int(run)             # index of the run (cli.py: `for run in range(N_RUNS)`): part of the sub-seeds
        self.config = experiment.config
        self.path_log = experiment.path_results / 'experiment.log'
        self.path_results = experiment.path_results
        self.model = Model(data=self.data, config=self.config['model'])
        self.ops = {}
        self.steps_per_operator()
        self.sampler = None
        self.samples = None
        self.sample_from_warm_up = None
        self.logger = experiment.logger

    def log_setup(self):
        c = self.config['mcmc']
        self.logger.info(self.model.get_setup_message())
        prec = c['PROPOSAL_PRECISION']
        inherit = ''
        if prec.get('inheritance') is not None:
            inherit = ('Pseudocounts for tuning the width of the proposal distribution for inheritance (beta): %s\n'
                       % prec['inheritance'])
        logging.info(
            '\nMCMC SETUP\n##########################################\n'
            f"MCMC with {c['N_STEPS']} steps and {c['N_SAMPLES']} samples\n"
            f"Warm-up: {c['WARM_UP']['N_WARM_UP_CHAINS']} chains exploring the parameter space in "
            f"{c['WARM_UP']['N_WARM_UP_STEPS']} steps\n"
            f"Pseudocounts for tuning the width of the proposal distribution for weights: {prec['weights']}\n"
            f"Pseudocounts for tuning the width of the proposal distribution for universal pressure (alpha): "
            f"{prec['universal']}\n{inherit}"
            f"Pseudocounts for tuning the width of the proposal distribution for areas (gamma): {prec['contact']}\n"
            f"Ratio of areal steps (growing, shrinking, swapping areas): {c['STEPS']['area']}\n"
            f"Ratio of weight steps (changing weights): {c['STEPS']['weights']}\n"
            f"Ratio of universal steps (changing alpha) : {c['STEPS']['universal']}\n"
            f"Ratio of inheritance steps (changing beta): {c['STEPS']['inheritance']}\n"
            f"Ratio of contact steps (changing gamma): {c['STEPS']['contact']}\n")

    def steps_per_operator(self):
        """mcmc_setup.py:70-95: area steps split 0.4 / 0.4 / 0.2 over shrink / grow / swap; with SAMPLE_SOURCE the
        parameter steps are the Gibbs operators, otherwise the Metropolis `alter_*` ones."""
        steps = self.config['mcmc']['STEPS']
        ops = {'shrink_zone': steps['area'] * 0.4, 'grow_zone': steps['area'] * 0.4, 'swap_zone': steps['area'] * 0.2,
               'gibbsish_sample_zones': steps['area'] * 0.0}
        kind = 'gibbs_sample' if self.model.sample_source else 'alter'
        if self.model.sample_source:
            ops['gibbs_sample_sources'] = steps['source']
        ops.update({f'{kind}_weights': steps['weights'], f'{kind}_p_global': steps['universal'],
                    f'{kind}_p_zones': steps['contact'], f'{kind}_p_families': steps['inheritance']})
        self.ops = ops

    def _sampler_kwargs(self, phase):
        """`phase` 0 = warm-up, 1 = sampling.  The config's SEED (an extension over the reference, which seeds nothing)
        is not handed to the samplers as it is: the device RNG is Philox keyed by (seed, chain) with a step counter that
        restarts at 0 in every sampler, so warm-up and sampling -- and every run of N_RUNS, every value of N_AREAS --
        would replay the same streams.  Each sampler instance gets its own sub-seed instead."""
        c = self.config['mcmc']
        kw = dict(data=self.data, model=self.model, operators=self.ops, var_proposal=c['PROPOSAL_PRECISION'],
                  p_grow_connected=c['P_GROW_CONNECTED'], initial_size=c['M_INITIAL'], logger=self.logger)
        if c.get('SEED') is not None:
            kw['seed'] = self.sub_seed(c['SEED'], self.run, self.config['model']['N_AREAS'], phase)
        return kw

    @staticmethod
    def sub_seed(seed, run, n_areas, phase):
        ss = np.random.SeedSequence([int(seed), int(run), int(n_areas), int(phase)])
        return int(ss.generate_state(1, dtype=np.uint64)[0] >> np.uint64(1))

    def sample(self, lh_per_area=True, initial_sample=None):
        c = self.config['mcmc']
        if initial_sample is None:
            initial_sample = self.sample_from_warm_up
        mc3 = c.get('MC3') or {}
        self.sampler = ZoneMCMC(n_chains=c.get('N_CHAINS', 1), initial_sample=initial_sample,
                                mc3=bool(mc3.get('ACTIVATE', False)), swap_period=mc3.get('SWAP_PERIOD'),
                                chain_swaps=mc3.get('N_SWAPS'), **self._sampler_kwargs(phase=1))
        self.sampler.generate_samples(c['N_STEPS'], c['N_SAMPLES'])
        if lh_per_area:
            contribution_per_area(self.sampler)
        self.samples = self.sampler.statistics

    def eval_ground_truth(self, lh_per_area=True):
        """mcmc_setup.py:121-172 (simulated data): likelihood / prior of the true parameters, per area too."""
        if self.config['model']['INHERITANCE']:
            weights = self.data.weights.copy()
        else:
            w = self.data.weights[:, :2]
            weights = w / w.sum(axis=-1, keepdims=True)
        p_global = self.data.p_universal[np.newaxis, ...]
        truth = Sample(zones=self.data.areas, weights=weights, p_global=p_global, p_zones=self.data.p_contact,
                       p_families=self.data.p_inheritance)
        s = self.samples
        s['true_zones'], s['true_weights'], s['true_p_global'] = self.data.areas, weights, p_global
        s['true_p_zones'], s['true_p_families'] = self.data.p_contact, self.data.p_inheritance
        s['true_ll'] = self.sampler.likelihood(truth, 0)
        s['true_prior'] = self.sampler.prior(truth, 0)
        s['true_families'] = self.data.families
        if lh_per_area:
            lhs, priors = [], []
            for z in range(len(self.data.areas)):
                single = Sample(zones=self.data.areas[np.newaxis, z], weights=weights, p_global=p_global,
                                p_zones=self.data.p_contact[np.newaxis, z], p_families=self.data.p_inheritance)
                lhs.append(self.sampler.likelihood(single, 0))
                priors.append(self.sampler.prior(single, 0))
            s['true_lh_single_zones'] = lhs
            s['true_prior_single_zones'] = priors
            s['true_posterior_single_zones'] = [a + b for a, b in zip(lhs, priors)]

    def warm_up(self):
        c = self.config['mcmc']
        warmup = ZoneMCMCWarmup(n_chains=c['WARM_UP']['N_WARM_UP_CHAINS'], **self._sampler_kwargs(phase=0))
        self.sample_from_warm_up = warmup.generate_samples(n_steps=0, n_samples=0, warm_up=True,
                                                           warm_up_steps=c['WARM_UP']['N_WARM_UP_STEPS'])
        warmup.close()

    def save_samples(self, run=1):
        """mcmc_setup.py:187-231: label matching, ranking, then `stats_<fi>_<run>.txt` / `areas_<fi>_<run>.txt`."""
        self.samples = match_areas(self.samples)
        self.samples = rank_areas(self.samples)
        file_info = self.config['results']['FILE_INFO']
        if file_info == "n":
            fi = 'n{n}'.format(n=self.config['model']['N_AREAS'])
        elif file_info == "s":
            fi = 's{s}a{a}'.format(s=self.config['simulation']['STRENGTH'], a=self.config['simulation']['AREA'])
        elif file_info == "i":
            fi = 'i{i}'.format(i=int(self.config['model']['INHERITANCE']))
        elif file_info == "p":
            fi = 'p{p}'.format(p=0 if self.config['model']['PRIOR']['universal'] == "uniform" else 1)
        else:
            raise ValueError("file_info must be 'n', 's', 'i' or 'p'")
        pth = self.path_results / fi
        gt_pth = pth / 'ground_truth'
        paths = {'parameters': pth / f'stats_{fi}_{run}.txt', 'areas': pth / f'areas_{fi}_{run}.txt',
                 'gt': gt_pth / 'stats.txt', 'gt_areas': gt_pth / 'areas.txt'}
        pth.mkdir(exist_ok=True)
        if self.data.is_simulated:
            self.eval_ground_truth()
            gt_pth.mkdir(exist_ok=True)
        samples2file(self.samples, self.data, self.config, paths)

    def log_statistics(self):
        self.sampler.print_statistics(self.samples)
