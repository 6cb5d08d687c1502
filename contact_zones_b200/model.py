"""Drop-in mirror of `sbayes/model.py` for the hot path: `Model`, `Likelihood`, `Prior` and the prior-term classes keep
the reference's constructor signatures, attributes, return types and error behaviour, but every number is computed by
the CUDA kernels behind the C ABI (`engine.Engine`): K1 (likelihood), K2 (size / Dirichlet priors), K3 (geo prior).
There is no host arithmetic fallback -- without libsbayes_b200.so and a GPU these calls raise."""
from enum import Enum

import numpy as np

from . import device as _dev

EPS = np.finfo(float).eps


def _mark_updated(sample, which, inheritance):
    """`Likelihood.everything_updated` (model.py:186-192) / `Prior.everything_updated` (:507-514)."""
    wc = getattr(sample, 'what_changed', None)
    if not wc:
        return
    w = wc[which]
    for k in ('zones', 'p_global', 'p_zones'):
        w[k].clear()
    w['weights'] = False
    if inheritance:
        w['p_families'].clear()


class Model(object):
    """`sbayes.model.Model` (model.py:14-66)."""

    def __init__(self, data, config):
        self.data = data
        self.config = config
        self.parse_attributes(config)
        self.likelihood = Likelihood(data=data, inheritance=self.inheritance)
        self.prior = Prior(data=data, inheritance=self.inheritance, prior_config=config['PRIOR'])

    def parse_attributes(self, config):
        self.n_zones = config['N_AREAS']
        self.min_size = config['MIN_M']
        self.max_size = config['MAX_M']
        self.inheritance = config['INHERITANCE']
        self.sample_source = config['SAMPLE_SOURCE']

    def __call__(self, sample, caching=True):
        return self.likelihood(sample, caching=caching) + self.prior(sample)

    def __copy__(self):
        return Model(self.data, self.config)

    def get_setup_message(self):
        msg = '\nModel\n##########################################\n'
        msg += f'Number of inferred areas: {self.n_zones}\n'
        msg += f'Areas have a minimum size of {self.min_size} and a maximum size of {self.max_size}\n'
        msg += f'Inheritance is considered for inference: {self.inheritance}\n'
        msg += self.prior.get_setup_message()
        return msg


class Likelihood(object):
    """`sbayes.model.Likelihood` (model.py:69-294).  Needs only `data.features`, `data.families`, `data.states` and
    `data.network` (test/test_model.py:76-80 builds such a stub)."""

    def __init__(self, data, inheritance):
        self.data = data
        self.features = data.features
        self.families = np.asarray(data.families, dtype=bool)
        self.inheritance = inheritance
        self.n_sites, self.n_features, self.n_categories = data.features.shape
        self.na_features = (np.sum(self.features, axis=-1) == 0)
        self.has_global = np.ones(self.n_sites, dtype=bool)
        self.has_family = np.any(self.families, axis=0) if self.families.size else np.zeros(self.n_sites, bool)
        self.has_zone = None
        self.has_components = None
        self.all_lh = None
        self.weights = None

    def reset_cache(self):
        """The device keeps no per-sample cache (a full K1 evaluation costs microseconds)."""
        self.has_zone = self.has_components = self.all_lh = self.weights = None

    def _engine(self, sample):
        return _dev.engine_for(self.data, np.asarray(sample.zones).shape[0], self.inheritance,
                               sample.source is not None)

    def __call__(self, sample, caching=True):
        """model.py:145-171 -> float (may be -inf, :181-182)."""
        eng = self._engine(sample)
        _dev.upload_sample(eng, 0, sample)
        ll, _, _ = eng.eval(0, 1, use_source=sample.source is not None)
        _mark_updated(sample, 'lh', self.inheritance)
        return float(ll[0])

    def everything_updated(self, sample):
        _mark_updated(sample, 'lh', self.inheritance)

    # per-cell arrays used by the reference's operators (zone_sampling.py:191-202) and by its tests
    def _components(self, sample):
        eng = self._engine(sample)
        _dev.upload_sample(eng, 0, sample)
        self.all_lh, self.weights = eng.components(0)
        return self.all_lh, self.weights

    def update_component_likelihoods(self, sample, caching=True):
        return self._components(sample)[0]

    def update_weights(self, sample):
        return self._components(sample)[1]

    def get_zone_assignment(self, sample):
        self.has_zone = np.any(sample.zones, axis=0)
        return self.has_zone

    def get_global_lh(self, sample):
        return self._components(sample)[0][:, :, 0]

    def get_zone_lh(self, sample):
        return self._components(sample)[0][:, :, 1]

    def get_family_lh(self, sample):
        if not self.inheritance:
            return None
        return self._components(sample)[0][:, :, 2]


def normalize_weights(weights, has_components):
    """model.py:436-452 as a host helper for callers that pass explicit arrays (tests)."""
    w = np.repeat(weights[np.newaxis, :, :], has_components.shape[0], axis=0) * has_components[:, np.newaxis, :]
    return w / w.sum(axis=2, keepdims=True)


def scale_counts(counts, scale_to, prior_inheritance=False):
    """util.py:569-586."""
    s = np.sum(counts, axis=-1)
    s = np.where(s == 0, EPS, s)
    factor = scale_to / s
    factor = np.where(factor < 1, factor, 1)
    return counts * factor[..., None]


# ---------------------------------------------------------------------------------------------------------------
# priors (model.py:455-1056)
# ---------------------------------------------------------------------------------------------------------------
def _prior_tables(prior_p_global, prior_p_families, inheritance):
    """Host-side constants of the Dirichlet priors in the form `sb_set_priors` takes (quirk Q10: the prior DENSITY
    uses alpha = prior.counts + 1, util.py:609,620, while the Gibbs operators use prior.counts, zone_sampling.py:344)."""
    t = {'p_global_type': 'uniform', 'p_families_type': 'uniform', 'gibbs_counts_global': None, 'alpha_global': None,
         'gibbs_counts_fam': None, 'alpha_fam': None}
    if prior_p_global is not None:
        t['gibbs_counts_global'] = np.asarray(prior_p_global.counts, dtype=float)
        if prior_p_global.prior_type is PGlobalPrior.TYPES.COUNTS:
            t['p_global_type'] = 'counts'
            t['alpha_global'] = t['gibbs_counts_global'] + 1.
    if inheritance and prior_p_families is not None:
        if prior_p_families.prior_type in (PFamiliesPrior.TYPES.UNIVERSAL, PFamiliesPrior.TYPES.COUNTS_AND_UNIVERSAL):
            # in the reference these two types fail at the first evaluation (prior_p_families_distr is never
            # initialised, model.py:726-748)
            raise NotImplementedError(f'PFamiliesPrior type {prior_p_families.prior_type.value!r} is not evaluated '
                                      'on the device')
        t['gibbs_counts_fam'] = np.asarray(prior_p_families.counts, dtype=float)
        if prior_p_families.prior_type is PFamiliesPrior.TYPES.COUNTS:
            t['p_families_type'] = 'counts'
            t['alpha_fam'] = t['gibbs_counts_fam'] + 1.
    return t


def _eval_prior(data, sample, inheritance, token, **prior_kwargs):
    """Upload `sample`, evaluate K2/K3 with exactly the prior terms in `prior_kwargs` switched on."""
    with_cost = prior_kwargs.get('geo_type', 'uniform') == 'cost_based'
    zones = np.asarray(sample.zones)
    if inheritance is None:        # a single prior term evaluated on its own: take the mixture width from the sample
        inheritance = np.asarray(sample.weights).shape[-1] == 3
    if with_cost and zones.shape[0] and np.any(zones.sum(axis=1) < 2):
        raise ValueError("Too few locations to compute distance.")           # model.py:1134-1135
    eng = _dev.engine_for(data, zones.shape[0], inheritance, False, with_cost=with_cost)
    if eng.prior_token is not token:
        eng.set_priors(**prior_kwargs)
        eng.prior_token = token
    _dev.upload_sample(eng, 0, sample)
    _, prior, geo = eng.eval(0, 1, use_source=False)
    eng.prior_token = token
    return float(prior[0]), float(geo[0])


class Prior(object):
    """`sbayes.model.Prior` (model.py:455-530): the joint log-prior, one K2 (+K3) launch."""

    def __init__(self, data, inheritance, prior_config):
        self.inheritance = inheritance
        self.data = data
        self.network = data.network
        self.config = prior_config
        self.size_prior = ZoneSizePrior(config=prior_config['area_size'], data=data)
        self.geo_prior = GeoPrior(config=prior_config['geo'], data=data)
        self.prior_weights = WeightsPrior(config=prior_config['weights'], data=data)
        self.prior_p_global = PGlobalPrior(config=prior_config['universal'], data=data)
        self.prior_p_zones = PZonesPrior(config=prior_config['contact'], data=data)
        if self.inheritance:
            self.prior_p_families = PFamiliesPrior(config=prior_config['inheritance'], data=data)
        self._token = object()

    def device_config(self):
        """Everything `Engine.set_priors` needs (also used by the samplers)."""
        kw = _prior_tables(self.prior_p_global, self.prior_p_families if self.inheritance else None, self.inheritance)
        kw['size_type'] = self.size_prior.prior_type.value
        kw['geo_type'] = self.geo_prior.prior_type.value
        kw['geo_scale'] = float(self.geo_prior.scale) if self.geo_prior.scale is not None else 1.0
        return kw

    def __call__(self, sample):
        if self.prior_p_zones.prior_type is not PZonesPrior.TYPES.UNIFORM:
            raise ValueError('Currently only uniform p_zones priors are supported.')      # model.py:826
        prior, _ = _eval_prior(self.data, sample, self.inheritance, self._token, **self.device_config())
        self.everything_updated(sample)
        return prior

    def everything_updated(self, sample):
        _mark_updated(sample, 'prior', self.inheritance)

    def get_setup_message(self):
        msg = self.geo_prior.get_setup_message()
        msg += self.size_prior.get_setup_message()
        msg += self.prior_weights.get_setup_message()
        msg += self.prior_p_global.get_setup_message()
        msg += self.prior_p_zones.get_setup_message()
        if self.inheritance:
            msg += self.prior_p_families.get_setup_message()
        return msg

    def __copy__(self):
        return Prior(data=self.data, inheritance=self.inheritance, prior_config=self.config)


class DirichletPrior(object):
    """model.py:533-565."""

    class TYPES(Enum):
        ...

    def __init__(self, config, data, initial_counts=1.):
        self.config = config
        self.data = data
        self.states = self.data.states
        self.initial_counts = initial_counts
        self.prior_type = None
        self.counts = None
        self.dirichlet = None
        self.cached = None
        self._token = object()
        self.parse_attributes(config)

    def parse_attributes(self, config):
        raise NotImplementedError()

    def is_outdated(self, sample):
        return self.cached is None

    def __call__(self, sample):
        raise NotImplementedError()

    def invalid_prior_message(self, s):
        name = self.__class__.__name__
        valid_types = ','.join([str(t.value) for t in self.TYPES])
        return f'Invalid prior type {s} for {name} (choose from [{valid_types}]).'


class PGlobalPrior(DirichletPrior):
    """model.py:567-630."""

    class TYPES(Enum):
        UNIFORM = 'uniform'
        COUNTS = 'counts'

    def parse_attributes(self, config):
        _, n_features, n_states = self.data.features.shape
        if config['type'] == 'uniform':
            self.prior_type = self.TYPES.UNIFORM
            self.counts = np.full(shape=(n_features, n_states), fill_value=self.initial_counts)
        elif config['type'] == 'counts':
            self.prior_type = self.TYPES.COUNTS
            if config['scale_counts'] is not None:      # mutates data in place like the reference (quirk Q10)
                self.data.prior_universal['counts'] = scale_counts(counts=self.data.prior_universal['counts'],
                                                                   scale_to=config['scale_counts'])
            self.counts = self.initial_counts + self.data.prior_universal['counts']
            self.dirichlet = [self.counts[f, np.asarray(self.states[f], bool)] + 1 for f in range(n_features)]
        else:
            raise ValueError(self.invalid_prior_message(config['type']))

    def __call__(self, sample):
        if self.prior_type is self.TYPES.UNIFORM:
            self.cached = 0
            return 0
        kw = _prior_tables(self, None, False)
        self.cached, _ = _eval_prior(self.data, sample, None, self._token, **kw)
        return self.cached

    def get_setup_message(self):
        msg = f'Universal-prior type: {self.prior_type.value}\n'
        if self.prior_type == self.TYPES.COUNTS:
            msg += f'\tCounts file: {self.config["file"]}\n'
        return msg


class PFamiliesPrior(DirichletPrior):
    """model.py:633-771."""

    class TYPES(Enum):
        UNIFORM = 'uniform'
        COUNTS = 'counts'
        UNIVERSAL = 'universal'
        COUNTS_AND_UNIVERSAL = 'counts_and_universal'

    def parse_attributes(self, config):
        if config['type'] == 'uniform':
            n_families, _ = self.data.families.shape
            _, n_features, n_states = self.data.features.shape
            self.prior_type = self.TYPES.UNIFORM
            self.counts = np.full(shape=(n_families, n_features, n_states), fill_value=self.initial_counts)
        elif config['type'] == 'universal':
            self.prior_type = self.TYPES.UNIVERSAL
            self.strength = config['scale_counts']
        elif config['type'] in ('counts', 'counts_and_universal'):
            self.prior_type = self.TYPES.COUNTS if config['type'] == 'counts' else self.TYPES.COUNTS_AND_UNIVERSAL
            if config['scale_counts'] is not None:
                self.data.prior_inheritance['counts'] = scale_counts(counts=self.data.prior_inheritance['counts'],
                                                                     scale_to=config['scale_counts'],
                                                                     prior_inheritance=True)
            if self.prior_type is self.TYPES.COUNTS:
                self.counts = self.initial_counts + self.data.prior_inheritance['counts']
                st = np.asarray(self.states, bool)
                self.dirichlet = [[self.counts[k, f, st[f]] + 1 for f in range(st.shape[0])]
                                  for k in range(self.counts.shape[0])]
            else:
                self.strength = config['scale_counts']
        else:
            raise ValueError(self.invalid_prior_message(config['type']))

    def __call__(self, sample):
        if self.prior_type is self.TYPES.UNIFORM:
            self.cached = 0.
            return 0.
        kw = _prior_tables(None, self, True)
        self.cached, _ = _eval_prior(self.data, sample, True, self._token, **kw)
        return self.cached

    def get_setup_message(self):
        msg = f'Prior on inheritance (beta): {self.prior_type.value}\n'
        if self.prior_type in [self.TYPES.COUNTS, self.TYPES.COUNTS_AND_UNIVERSAL]:
            msg += '\tCounts files:\n'
            for fam, path in self.config['files'].items():
                msg += f'\t\t{fam}: {path}\n'
        if self.prior_type in [self.TYPES.UNIVERSAL, self.TYPES.COUNTS_AND_UNIVERSAL]:
            msg += f'\tUniversal hyperprior strength: {self.strength}\n'
        return msg


class PZonesPrior(DirichletPrior):
    """model.py:774-848: only `uniform` evaluates (the `universal` type raises in the reference, :826)."""

    class TYPES(Enum):
        UNIFORM = 'uniform'
        UNIVERSAL = 'universal'

    def parse_attributes(self, config):
        _, n_features, n_states = self.data.features.shape
        if config['type'] == 'uniform':
            self.prior_type = self.TYPES.UNIFORM
            self.counts = np.full(shape=(n_features, n_states), fill_value=self.initial_counts)
        elif config['type'] == 'universal':
            self.prior_type = self.TYPES.UNIVERSAL
            self.strength = config['scale_counts']
            self.states = self.data.state_names['internal']
        else:
            raise ValueError(self.invalid_prior_message(config['type']))

    def __call__(self, sample):
        if self.prior_type == self.TYPES.UNIFORM:
            self.cached = 0.
            return 0.
        raise ValueError('Currently only uniform p_zones priors are supported.')

    def get_setup_message(self):
        return f'Prior on contact (gamma): {self.prior_type.value}\n'


class WeightsPrior(DirichletPrior):
    """model.py:851-890."""

    class TYPES(Enum):
        UNIFORM = 'uniform'

    def parse_attributes(self, config):
        _, n_features, n_states = self.data.features.shape
        if config['type'] == 'uniform':
            self.prior_type = self.TYPES.UNIFORM
            self.counts = np.full(shape=(n_features, n_states), fill_value=self.initial_counts)
        else:
            raise ValueError(self.invalid_prior_message(config['type']))

    def __call__(self, sample):
        self.cached = 0.
        return 0.

    def get_setup_message(self):
        return f'Prior on weights: {self.prior_type.value}\n'


class ZoneSizePrior(object):
    """model.py:893-976; evaluated by K2 (`log_binom` via lgamma, util.py:1202-1217)."""

    class TYPES(Enum):
        UNIFORM_AREA = 'none'
        UNIFORM_SIZE = 'uniform'
        QUADRATIC_SIZE = 'quadratic'

    def __init__(self, config, data, initial_counts=1.):
        self.config = config
        self.data = data
        self.states = self.data.states
        self.initial_counts = initial_counts
        self.prior_type = None
        self.counts = None
        self.dirichlet = None
        self.cached = None
        self._token = object()
        self.parse_attributes(config)

    def invalid_prior_message(self, s):
        valid_types = ','.join([str(t.value) for t in self.TYPES])
        return f'Invalid prior type {s} for size prior (choose from [{valid_types}]).'

    def parse_attributes(self, config):
        try:
            self.prior_type = self.TYPES(config['type'])
        except ValueError:
            raise ValueError(self.invalid_prior_message(config['type']))

    def __call__(self, sample):
        if self.prior_type is self.TYPES.UNIFORM_AREA:
            self.cached = 0.
        else:
            self.cached, _ = _eval_prior(self.data, sample, None, self._token, size_type=self.prior_type.value)
        return self.cached

    def get_setup_message(self):
        return f'Prior on area size: {self.prior_type.value}\n'


class GeoPrior(object):
    """model.py:979-1056; `cost_based` is K3 (Prim over the last zone's members, quirk Q1)."""

    class TYPES(Enum):
        UNIFORM = 'uniform'
        COST_BASED = 'cost_based'

    def __init__(self, config, data, initial_counts=1.):
        self.config = config
        self.data = data
        self.states = self.data.states
        self.initial_counts = initial_counts
        self.prior_type = None
        self.cost_matrix = None
        self.scale = None
        self.cached = None
        self._token = object()
        self.parse_attributes(config)

    def parse_attributes(self, config):
        if config['type'] == 'uniform':
            self.prior_type = self.TYPES.UNIFORM
        elif config['type'] == 'cost_based':
            self.prior_type = self.TYPES.COST_BASED
            self.cost_matrix = self.data.geo_prior['cost_matrix']         # None: implicit (site coordinates), f3
            self.scale = config['scale']
        else:
            raise ValueError('Geo prior not supported')

    def __call__(self, sample):
        if self.prior_type is self.TYPES.UNIFORM:
            self.cached = 0.
        else:
            _, self.cached = _eval_prior(self.data, sample, None, self._token, geo_type='cost_based',
                                         geo_scale=float(self.scale))
        return self.cached

    def invalid_prior_message(self, s):
        valid_types = ','.join([str(t.value) for t in self.TYPES])
        return f'Invalid prior type {s} for geo-prior (choose from [{valid_types}]).'

    def get_setup_message(self):
        msg = f'Geo-prior: {self.prior_type.value}\n'
        if self.prior_type == self.TYPES.COST_BASED:
            msg += f'\tScale: {self.scale}\n'
        return msg
