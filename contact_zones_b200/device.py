"""Engine pool: one C-ABI context per (data object, model shape) so that the reference-shaped objects in `model.py`
(`Likelihood`, `Prior`, ...) can evaluate arbitrary host `Sample`s without re-uploading the immutable data on every
call.  The reference keeps the equivalent caches inside `Likelihood` (`model.py:118-131`)."""
import os

import numpy as np

from .engine import Engine

_POOL = {}


def default_device():
    """One process per GPU: LOCAL_RANK (torchrun) selects the device, else device 0."""
    return int(os.environ.get('LOCAL_RANK', '0'))


def cost_matrix_of(data):
    gp = getattr(data, 'geo_prior', None)
    if isinstance(gp, dict):
        return gp.get('cost_matrix')
    return None


def engine_for(data, n_zones, inheritance, sample_source, n_chains=1, with_cost=False, warmup=False, device=None,
               chain_offset=0, pooled=True):
    """Context for evaluating states with `n_zones` zones over `data`.  Pooled contexts are shared (evaluation API);
    samplers ask for their own (`pooled=False`) because they hold chain state across calls."""
    device = default_device() if device is None else int(device)
    key = (id(data), int(n_zones), bool(inheritance), bool(sample_source), int(n_chains), bool(with_cost), bool(warmup),
           device, int(chain_offset))
    if pooled and key in _POOL:
        return _POOL[key][0]
    cost = cost_matrix_of(data) if with_cost else None
    locations = None
    if with_cost and cost is None:
        # no dense matrix: the implicit Euclidean cost over the site coordinates, if the data object carries them
        # (`Data.load_geo_cost_matrix` with DENSE_DISTANCES = false)
        gp = getattr(data, 'geo_prior', None)
        locations = gp.get('locations') if isinstance(gp, dict) else None
        if locations is None:
            raise KeyError('cost_matrix')        # the reference fails the same way at model.py:1009
    network = getattr(data, 'network', None)
    adj = network['adj_mat'] if network else None
    states = getattr(data, 'states', None)       # test/test_model.py:76-80 passes a stub with features + families only
    if states is None:
        states = np.ones(np.asarray(data.features).shape[1:], dtype=bool)
    feats = getattr(data, 'features_idx', None)        # uint8 state index straight from the encoder (data.py)
    if feats is None:
        feats = data.features
    eng = Engine(feats, states, data.families if inheritance else None, adj, cost,
                 n_zones=n_zones, n_chains=n_chains, inheritance=inheritance, sample_source=sample_source, warmup=warmup,
                 device=device, chain_offset=chain_offset, locations=locations)
    eng.prior_token = None
    if pooled:
        _POOL[key] = (eng, data)             # keeping `data` alive pins id(data)
    return eng


def clear_pool():
    for eng, _ in _POOL.values():
        eng.close()
    _POOL.clear()


def upload_sample(eng, chain, sample):
    """Host `Sample` (zone_sampling.py:49-75) -> device image of `chain`."""
    src = sample.source if eng.sample_source else None
    if eng.sample_source and src is None:
        # a source-less Sample on a SAMPLE_SOURCE context: component 0 (universal) is always available
        src = np.zeros((eng.N, eng.F), np.uint8)
    eng.set_state(chain, sample.zones, sample.weights, sample.p_global, sample.p_zones, sample.p_families, src)
