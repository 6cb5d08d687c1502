"""Callers and file formats directly around the hot path (SURVEY.md 8f rows 1-2, 8b "CLI + files"):
`contribution_per_area` (postprocessing.py:271-313) as ONE batched device evaluation over (sample, zone) pairs,
`match_areas` (:205-268) without the O(Z! * N * Z) inner product, `rank_areas` (:316-365) and the `stats_*.txt` /
`areas_*.txt` writers (util.py:647-907) with the reference's column names and order."""
import csv
from itertools import permutations

import numpy as np

from . import device as _dev


def contribution_per_area(mcmc_sampler, batch=256):
    """Likelihood and prior of every zone ALONE for every logged sample.  The reference builds N_SAMPLES * Z
    single-zone `Sample`s and evaluates them one by one (`postprocessing.py:287-303`); here they fill the chain slots
    of a Z=1 context and each batch is one K1 + K2/K3 launch."""
    stats = mcmc_sampler.statistics
    model = mcmc_sampler.model
    n_samples = len(stats['sample_zones'])
    pairs = [(s, z) for s in range(n_samples) for z in range(len(stats['sample_zones'][s]))]
    lh = [[None] * len(stats['sample_zones'][s]) for s in range(n_samples)]
    pr = [[None] * len(stats['sample_zones'][s]) for s in range(n_samples)]
    if pairs:
        kw = model.prior.device_config()
        batch = min(batch, len(pairs))
        eng = _dev.engine_for(mcmc_sampler.data, 1, model.inheritance, False, n_chains=batch,
                              with_cost=kw['geo_type'] == 'cost_based')
        eng.set_priors(**kw)
        eng.prior_token = None
        inh = model.inheritance
        for b0 in range(0, len(pairs), batch):
            chunk = pairs[b0:b0 + batch]
            zones = np.stack([np.asarray(stats['sample_zones'][s])[np.newaxis, z] for s, z in chunk])
            if kw['geo_type'] == 'cost_based' and (zones.reshape(len(chunk), -1).sum(axis=1) < 2).any():
                raise ValueError("Too few locations to compute distance.")
            # ONE bulk upload per batch (sb_set_states): the (sample, zone) pairs fill consecutive chain slots
            eng.set_states(0, zones, np.stack([stats['sample_weights'][s] for s, _ in chunk]),
                           np.stack([np.asarray(stats['sample_p_global'][s]) for s, _ in chunk]),
                           np.stack([np.asarray(stats['sample_p_zones'][s])[np.newaxis, z] for s, z in chunk]),
                           np.stack([np.asarray(stats['sample_p_families'][s]) for s, _ in chunk]) if inh else None, None)
            ll, prior, _ = eng.eval(0, len(chunk), use_source=False)
            for slot, (s, z) in enumerate(chunk):
                lh[s][z], pr[s][z] = float(ll[slot]), float(prior[slot])
    stats['sample_lh_single_zones'] = lh
    stats['sample_prior_single_zones'] = pr
    stats['sample_posterior_single_zones'] = [[a + b for a, b in zip(l, p)] for l, p in zip(lh, pr)]
    mcmc_sampler.statistics = stats


def match_areas(samples):
    """Label switching: relabel the zones of every sample so that they agree best with the running sum of the
    previous (relabelled) samples.  Same result as the reference, which scores all Z! permutations against the
    (N, Z) arrays (`postprocessing.py:219-236`): the score of permutation p is sum_i M[i, p_i] with the Z x Z overlap
    matrix M = s_sum^T s, so the N-sized work is done once per sample; ties resolve to the first permutation in
    `itertools.permutations` order, like `max()`."""
    zones = np.array([np.array(s) for s in samples['sample_zones']])        # (n_samples, Z, N)
    n_samples, n_areas, n_sites = zones.shape
    perm = np.array(list(permutations(range(n_areas))), dtype=int).reshape(-1, n_areas)
    rows = np.arange(n_areas)
    s_sum = np.zeros((n_areas, n_sites))
    matching = []
    for s in zones:
        overlap = s_sum @ s.T.astype(float)                                  # (Z, Z) integer-valued
        scores = overlap[rows[None, :], perm].sum(axis=1)
        best = perm[int(np.argmax(scores))]
        matching.append(list(best))
        s_sum += s[best]
    print("Matching areas ...")
    samples['sample_zones'] = [np.asarray(samples['sample_zones'][s])[matching[s]] for s in range(n_samples)]
    samples['sample_p_zones'] = [np.asarray(p)[matching[s]] for s, p in enumerate(samples['sample_p_zones'])]
    for key in ('sample_lh_single_zones', 'sample_prior_single_zones', 'sample_posterior_single_zones'):
        samples[key] = [[row[i] for i in matching[s]] for s, row in enumerate(samples[key])]
    return samples


def rank_areas(samples):
    """postprocessing.py:316-365: order zones by their mean single-zone posterior, best first."""
    post = np.asarray(samples['sample_posterior_single_zones'])
    ranked = np.argsort(-np.mean(post, axis=0))
    print("Ranking areas ...")
    samples['sample_zones'] = [np.asarray(z)[ranked] for z in samples['sample_zones']]
    samples['sample_p_zones'] = [np.asarray(p)[ranked] for p in samples['sample_p_zones']]
    for key in ('sample_lh_single_zones', 'sample_prior_single_zones', 'sample_posterior_single_zones'):
        samples[key] = [[row[r] for r in ranked] for row in samples[key]]
    return samples


# ---- writers (util.py:64-80, 647-907) ---------------------------------------------------------------------------
def encode_area(area):
    return ''.join(np.asarray(area).astype(int).astype(str))


def decode_area(area_str):
    return np.array(list(area_str)).astype(int).astype(bool)


def format_area_columns(areas):
    return '\t'.join(encode_area(a) for a in areas)


def parse_area_columns(areas_encoded):
    return np.array([decode_area(a) for a in areas_encoded.split('\t')])


def _parameter_columns(data, weights, p_global, p_zones, p_families, n_areas, inheritance, beta=None):
    """(name, value) pairs in the reference's order: weights, alpha, gamma, beta."""
    feats = data.feature_names['external']
    states = data.state_names['external']
    out = []
    beta = inheritance if beta is None else beta
    comp = ['universal', 'contact'] + (['inheritance'] if inheritance else [])
    for f, fname in enumerate(feats):
        for c, cname in enumerate(comp):
            out.append((f'w_{cname}_{fname}', weights[f][c]))
    for f, fname in enumerate(feats):
        for i, st in enumerate(states[f]):
            out.append((f'alpha_{fname}_{st}', p_global[0][f][states[f].index(st)]))
    for a in range(n_areas):
        for f, fname in enumerate(feats):
            for st in states[f]:
                out.append((f'gamma_a{a + 1}_{fname}_{st}', p_zones[a][f][states[f].index(st)]))
    if beta:
        for k, famname in enumerate(data.family_names['external']):
            for f, fname in enumerate(feats):
                for st in states[f]:
                    out.append((f'beta_{famname}_{fname}_{st}', p_families[k][f][states[f].index(st)]))
    return out


def collect_row_for_writing(s, samples, data, config, steps_per_sample):
    """util.py:744-845."""
    row = {'Sample': int(s * steps_per_sample),
           'posterior': samples['sample_prior'][s] + samples['sample_likelihood'][s],
           'likelihood': samples['sample_likelihood'][s], 'prior': samples['sample_prior'][s]}
    for i, area in enumerate(samples['sample_zones'][s]):
        row[f'size_a{i}'] = np.count_nonzero(area)
    inh = config['model']['INHERITANCE']
    row.update(_parameter_columns(data, samples['sample_weights'][s], samples['sample_p_global'][s],
                                  samples['sample_p_zones'][s], samples['sample_p_families'][s],
                                  config['model']['N_AREAS'], inh))
    if data.is_simulated:
        sample_z = np.any(samples['sample_zones'][s], axis=0)
        true_z = np.any(samples['true_zones'], axis=0)
        inter = np.minimum(sample_z, true_z)
        row['recall'] = np.sum(inter, axis=0) / np.sum(true_z)
        row['precision'] = np.sum(inter, axis=0) / np.sum(sample_z, axis=0)
    if 'sample_lh_single_zones' in samples:
        for a in range(config['model']['N_AREAS']):
            row[f'lh_a{a + 1}'] = samples['sample_lh_single_zones'][s][a]
            row[f'prior_a{a + 1}'] = samples['sample_prior_single_zones'][s][a]
            row[f'post_a{a + 1}'] = samples['sample_posterior_single_zones'][s][a]
    return row, list(row)


def collect_gt_for_writing(samples, data, config):
    """util.py:655-733."""
    gt = {'posterior': samples['true_prior'] + samples['true_ll'], 'likelihood': samples['true_ll'],
          'prior': samples['true_prior']}
    # the reference keys the weight columns on the model and the beta columns on the simulation config
    cols = _parameter_columns(data, samples['true_weights'], samples['true_p_global'], samples['true_p_zones'],
                              samples['true_p_families'], len(data.areas), config['model']['INHERITANCE'],
                              beta=config['simulation']['INHERITANCE'])
    gt.update(cols)
    if 'true_lh_single_zones' in samples:
        for a in range(len(data.areas)):
            gt[f'lh_a{a + 1}'] = samples['true_lh_single_zones'][a]
            gt[f'prior_a{a + 1}'] = samples['true_prior_single_zones'][a]
            gt[f'post_a{a + 1}'] = samples['true_posterior_single_zones'][a]
    return gt, list(gt)


def samples2file(samples, data, config, paths):
    """util.py:848-907: tab-separated `stats` file (one row per sample) and `areas` file (one line of Z bit-strings
    per sample); for simulated data also the ground-truth pair."""
    print("Writing results to file ...")
    if data.is_simulated:
        with open(paths['gt'], 'w', newline='') as fh:
            gt, names = collect_gt_for_writing(samples=samples, data=data, config=config)
            w = csv.DictWriter(fh, fieldnames=names, delimiter='\t')
            w.writeheader()
            w.writerow(gt)
        with open(paths['gt_areas'], 'w', newline='') as fh:
            fh.write(format_area_columns(samples['true_zones']))
    steps_per_sample = float(config['mcmc']['N_STEPS'] / config['mcmc']['N_SAMPLES'])
    with open(paths['parameters'], 'w', newline='') as fh:
        w = None
        for s in range(len(samples['sample_zones'])):
            row, names = collect_row_for_writing(s, samples, data, config, steps_per_sample)
            if w is None:
                w = csv.DictWriter(fh, fieldnames=names, delimiter='\t')
                w.writeheader()
            w.writerow(row)
    with open(paths['areas'], 'w', newline='') as fh:
        for s in range(len(samples['sample_zones'])):
            fh.write(format_area_columns(samples['sample_zones'][s]) + '\n')
