"""TEST INFRASTRUCTURE ONLY -- CPU (numpy) restatement of the reference's MCMC step loop.

`OracleSampler` restates `MCMCGenerative.step` (mcmc_generative.py:282-351) with the operators of `ZoneMCMC`
(zone_sampling.py:180-933) and the cached posterior evaluation of `Likelihood` / `Prior`
(model.py:145-294, 484-505).  It keeps the reference's cost structure on purpose -- a deep copy of the whole
state per copying operator (zone_sampling.py:95-115), dense (N,F,C) float64 arrays, the dirty-flag driven
partial recomputation of the component likelihoods, three full source-posterior passes per zone move -- because
it is (a) the timed CPU baseline of `bench.py` on the GPU box, where `/root/reference` does not exist, and
(b) the distributional checker for the CUDA sampler in `tests/`.  The product never imports it.

Pinned: `oracle/make_golden.py --runs` runs this sampler and the imported reference sampler on the same data
in the build container and stores the reference's posterior summaries in `tests/golden/run_*.npz`;
`tests/test_sampler_oracle.py` checks this port against them within Monte-Carlo error, and
`oracle/time_vs_reference.py` records that its step rate matches the reference's (DESIGN.md "CPU baseline").

RNG: numpy `Generator` (the reference interleaves stdlib `random` and `np.random` global state, quirk Q13, so
trajectories are not comparable bit for bit anyway).
"""
import math

import numpy as np

import sbayes_oracle as orc

Q_GIBBS, Q_BACK_GIBBS = -np.inf, 0.0          # mcmc_generative.py:27-31
Q_REJECT, Q_BACK_REJECT = 0.0, -np.inf


class State:
    """The arrays of a `Sample` (zone_sampling.py:49-75) plus its dirty flags (`what_changed`, :84-92)."""
    __slots__ = ('zones', 'weights', 'p_global', 'p_zones', 'p_families', 'source', 'dirty')

    def __init__(self, zones, weights, p_global, p_zones, p_families, source):
        self.zones, self.weights, self.p_global = zones, weights, p_global
        self.p_zones, self.p_families, self.source = p_zones, p_families, source
        self.dirty = None
        self.everything_changed()

    def everything_changed(self):
        self.dirty = {k: {'zones': 'all', 'weights': True, 'p_global': 'all', 'p_zones': 'all', 'p_families': 'all'}
                      for k in ('lh', 'prior')}

    def copy(self):
        """Sample.copy, zone_sampling.py:95-115: every array is copied on every copying operator."""
        new = State(self.zones.copy(), self.weights.copy(), self.p_global.copy(), self.p_zones.copy(),
                    None if self.p_families is None else self.p_families.copy(),
                    None if self.source is None else self.source.copy())
        new.dirty = {k: {n: (set(v) if isinstance(v, set) else v) for n, v in d.items()}
                     for k, d in self.dirty.items()}
        return new

    def mark(self, what, idx=None):
        for kind in ('lh', 'prior'):
            d = self.dirty[kind]
            if what == 'weights':
                d['weights'] = True
            elif d[what] != 'all':
                d[what].add(idx)


def _is_dirty(v):
    return v == 'all' or (isinstance(v, set) and len(v) > 0) or v is True


class CachedLikelihood:
    """Likelihood with the per-chain caches of model.py:105-294."""

    def __init__(self, features, families, inheritance):
        self.features = features
        self.families = np.asarray(families, dtype=bool)
        self.inheritance = inheritance
        self.N, self.F, self.S = features.shape
        self.na = features.sum(axis=-1) == 0
        self.has_family = self.families.any(axis=0) if inheritance else None
        self.reset()

    def reset(self):
        self.lh_global = self.lh_zone = self.lh_fam = self.all_lh = None
        self.has_zone = self.has = self.wnorm = None

    def components(self, st, caching=True):
        """update_component_likelihoods, model.py:230-249 (partial recomputation :297-433)."""
        d = st.dirty['lh']
        f, N, F = self.features, self.N, self.F
        if self.lh_global is None or _is_dirty(d['p_global']):
            if self.lh_global is None:
                self.lh_global = np.ones((N, F))
            idx = range(F) if (d['p_global'] == 'all') else d['p_global']
            for i in idx:
                self.lh_global[:, i] = f[:, i, :].dot(st.p_global[0, i, :])
        if self.inheritance and (self.lh_fam is None or _is_dirty(d['p_families'])):
            if self.lh_fam is None:
                self.lh_fam = np.zeros((N, F))
            if d['p_families'] == 'all':
                idx = ((k, i) for k in range(len(self.families)) for i in range(F))
            else:
                idx = d['p_families']
            for k, i in idx:
                m = self.families[k]
                self.lh_fam[m, i] = f[m, i, :].dot(st.p_families[k, i, :])
        if self.lh_zone is None or _is_dirty(d['zones']) or _is_dirty(d['p_zones']):
            if self.lh_zone is None:
                self.lh_zone = np.zeros((N, F))
            if d['p_zones'] == 'all' or d['zones'] == 'all':
                zs = range(len(st.zones))
            else:
                zs = set(d['zones']) | {z for z, _ in d['p_zones']}
            for z in zs:
                m = st.zones[z]
                self.lh_zone[m, :] = np.einsum('ijk,jk->ij', f[m], st.p_zones[z])
        if (not caching or self.all_lh is None or _is_dirty(d['zones']) or _is_dirty(d['p_global'])
                or _is_dirty(d['p_zones']) or _is_dirty(d['p_families'])):
            comps = [self.lh_global, self.lh_zone] + ([self.lh_fam] if self.inheritance else [])
            self.all_lh = np.array(comps).transpose((1, 2, 0))
            self.all_lh[self.na] = 1.
        return self.all_lh

    def zone_assignment(self, st):
        if self.has_zone is None or _is_dirty(st.dirty['lh']['zones']):
            self.has_zone = np.any(st.zones, axis=0)
        return self.has_zone

    def weights(self, st):
        """update_weights, model.py:257-294."""
        d = st.dirty['lh']
        self.zone_assignment(st)
        if self.has is None or _is_dirty(d['zones']):
            cols = [np.ones(self.N, dtype=bool), self.has_zone] + ([self.has_family] if self.inheritance else [])
            self.has = np.array(cols).T
        if self.wnorm is None or d['weights'] or _is_dirty(d['zones']):
            self.wnorm = orc.normalized_weights(st.weights, self.has)
        return self.wnorm

    def __call__(self, st, caching=True):
        if not caching:
            self.reset()
        with np.errstate(divide='ignore', invalid='ignore'):
            all_lh = self.components(st)
            w = self.weights(st)
            ll = orc.combine(all_lh, w, st.source)
        st.dirty['lh'] = {'zones': set(), 'weights': False, 'p_global': set(), 'p_zones': set(), 'p_families': set()}
        return ll


class CachedPrior:
    """Prior.__call__ (model.py:484-505) with the per-term caches (:932-972, 1018-1043, 596-623, 698-755)."""

    def __init__(self, states, inheritance, prior_cfg, cost_mat=None, alpha_global=None, alpha_families=None):
        self.states = states
        self.inheritance = inheritance
        self.cfg = prior_cfg
        self.cost = cost_mat
        self.alpha_g, self.alpha_f = alpha_global, alpha_families
        self.size = self.geo = None
        self.pg = self.pf = None

    def __call__(self, st):
        d = st.dirty['prior']
        cfg = self.cfg
        if self.size is None or _is_dirty(d['zones']):
            self.size = orc.size_prior(st.zones, cfg['area_size']['type'])
            self.geo = (orc.geo_prior_cost_based(st.zones, self.cost, cfg['geo']['scale'])
                        if cfg['geo']['type'] == 'cost_based' else 0.)
        total = self.size + self.geo
        if cfg['universal']['type'] == 'counts':
            F = st.p_global.shape[1]
            if self.pg is None:
                self.pg = np.zeros(F)
                idx = range(F)
            else:
                idx = range(F) if d['p_global'] == 'all' else d['p_global']
            for i in idx:
                s = self.states[i]
                self.pg[i] = orc.dirichlet_logpdf(st.p_global[0, i, s], self.alpha_g[i, s])
            total += self.pg.sum()
        if self.inheritance and cfg['inheritance']['type'] == 'counts':
            K, F = st.p_families.shape[:2]
            if self.pf is None:
                self.pf = np.zeros((K, F))
                idx = ((k, i) for k in range(K) for i in range(F))
            else:
                idx = ((k, i) for k in range(K) for i in range(F)) if d['p_families'] == 'all' else d['p_families']
            for k, i in idx:
                s = self.states[i]
                self.pf[k, i] = orc.dirichlet_logpdf(st.p_families[k, i, s], self.alpha_f[k, i, s])
            total += self.pf.sum()
        st.dirty['prior'] = {'zones': set(), 'weights': False, 'p_global': set(), 'p_zones': set(),
                             'p_families': set()}
        return float(total)


def sample_categorical(rng, p):
    """preprocessing.py:321-348 with binary encoding: first c with u < cumsum(p)[c] (argmax of an all-False row = 0)."""
    cdf = np.cumsum(p, axis=-1)
    u = rng.random(p.shape[:-1] + (1,))
    return np.eye(p.shape[-1], dtype=bool)[np.argmax(u < cdf, axis=-1)]


class OracleSampler:
    """One chain of ZoneMCMC.  `operators`: {name: weight} as built by MCMC.steps_per_operator (mcmc_setup.py:70-95)."""

    def __init__(self, features, states, families, adj_mat, inheritance, sample_source, prior_cfg, operators,
                 precision, min_size, max_size, p_grow_connected, state, seed=0, cost_mat=None,
                 gibbs_counts_global=None, alpha_global=None, gibbs_counts_fam=None, alpha_fam=None, warmup=False):
        self.features = np.asarray(features, dtype=bool)
        self.states = np.asarray(states, dtype=bool)
        self.families = np.asarray(families, dtype=bool)
        self.adj = adj_mat
        self.inheritance, self.sample_source, self.warmup = inheritance, sample_source, warmup
        self.N, self.F, self.S = self.features.shape
        self.Z = state['zones'].shape[0]
        self.K = self.families.shape[0] if inheritance else 0
        self.C = 3 if inheritance else 2
        self.precision = precision
        self.min_size, self.max_size, self.rho = min_size, max_size, p_grow_connected
        self.rng = np.random.default_rng(seed)
        self.lik = CachedLikelihood(self.features, self.families, inheritance)
        self.pri = CachedPrior(self.states, inheritance, prior_cfg, cost_mat, alpha_global, alpha_fam)
        self.gibbs_g = np.ones((self.F, self.S)) if gibbs_counts_global is None else gibbs_counts_global
        self.gibbs_f = np.ones((self.K, self.F, self.S)) if gibbs_counts_fam is None else gibbs_counts_fam
        self.op_names = [k for k, v in operators.items() if v > 0]
        p = np.array([operators[k] for k in self.op_names], dtype=float)
        self.op_p = p / p.sum()
        self.st = State(np.array(state['zones'], dtype=bool), np.array(state['weights'], dtype=float),
                        np.array(state['p_global'], dtype=float).reshape(1, self.F, self.S),
                        np.array(state['p_zones'], dtype=float),
                        None if not inheritance else np.array(state['p_families'], dtype=float),
                        None if not sample_source else np.array(state['source'], dtype=bool))
        self.ll = self.lik(self.st)
        self.prior = self.pri(self.st)
        self.accepted = {k: 0 for k in self.op_names}
        self.rejected = {k: 0 for k in self.op_names}
        self.n_steps = 0

    # -- MCMCGenerative.step, mcmc_generative.py:282-329 -------------------------------------------
    def step(self):
        name = self.op_names[self.rng.choice(len(self.op_names), p=self.op_p)]
        cand, log_q, log_q_back = getattr(self, name)(self.st)
        ll = self.lik(cand)            # evaluated even for forced rejects (quirk Q8)
        prior = self.pri(cand)
        if log_q_back == -np.inf:
            accept = False
        elif log_q == -np.inf:
            accept = True
        else:
            mh = (ll - self.ll) - (log_q - log_q_back) + (prior - self.prior)
            accept = math.log(self.rng.random()) < mh
        if accept:
            self.st, self.ll, self.prior = cand, ll, prior
            self.accepted[name] += 1
        else:
            self.rejected[name] += 1
        self.n_steps += 1

    def run(self, n_steps):
        for _ in range(n_steps):
            self.step()

    # -- sources (zone_sampling.py:180-220) ----------------------------------------------------------
    def _source_posterior(self, st):
        with np.errstate(divide='ignore', invalid='ignore'):
            lh = self.lik.components(st, caching=False)
            w = self.lik.weights(st)
            return orc.normalize(lh * w, axis=-1)

    def gibbs_sample_sources(self, st, as_gibbs=True):
        post = self._source_posterior(st)
        st.source = sample_categorical(self.rng, post)
        if as_gibbs:
            return st, Q_GIBBS, Q_BACK_GIBBS
        return st, float(np.sum(np.log(post.ravel()[np.where(st.source.ravel())]))), 0.

    def _log_q_back_sources(self, st):
        post = self._source_posterior(st)
        return float(np.sum(np.log(post.ravel()[np.where(st.source.ravel())])))

    # -- Gibbs operators (zone_sampling.py:222-406) ---------------------------------------------------
    def gibbs_sample_weights(self, st):
        new = st.copy()
        st.mark('weights')
        new.mark('weights')
        if not self.inheritance:
            counts = np.sum(st.source[self.lik.zone_assignment(st)], axis=0)
            for i in range(self.F):
                new.weights[i, :] = self.rng.dirichlet(1 + counts[i])
            return new, Q_GIBBS, Q_BACK_GIBBS
        w, w_new = st.weights, new.weights
        if self.rng.random() < 0.5:        # fixed = 'inheritance' -> resample contact vs universal
            counts = np.sum(st.source[self.lik.zone_assignment(st)], axis=0)
            col = 1
        else:                              # fixed = 'contact' -> resample inheritance vs universal
            counts = np.sum(st.source[self.lik.has_family], axis=0)
            col = 2
        a = self.rng.beta(1 + counts[:, col], 1 + counts[:, 0])
        w_new[:, col] = a * w[:, 0] / (1 - a)
        w_new = orc.normalize(w_new, axis=-1)
        # the reference also evaluates a per-feature MH ratio here (two (N,F,C) log passes, :306-325) and then
        # overwrites its decision (:327, quirk Q2); the cost is kept, the decision is not used
        with np.errstate(divide='ignore', invalid='ignore'):
            wn_old = self.lik.weights(st)
            np.sum(np.log(np.sum(st.source * wn_old, axis=-1)), axis=0)
            has = self.lik.has
            wn_new = orc.normalized_weights(w_new, has)
            np.sum(np.log(np.sum(new.source * wn_new, axis=-1)), axis=0)
        new.weights = w_new
        return new, Q_GIBBS, Q_BACK_GIBBS

    def _dirichlet_rows(self, table, counts, prior_counts, feats):
        for i in feats:
            s = self.states[i]
            table[i, s] = self.rng.dirichlet(prior_counts[i, s] + counts[i, s])

    def gibbs_sample_p_global(self, st, frac=0.4):
        subset = self.rng.random(self.F) < frac
        feats = np.flatnonzero(subset)
        f = st.source[:, subset, 0, np.newaxis] * self.features[:, subset, :]
        for j, i in enumerate(feats):
            s = self.states[i]
            st.p_global[0, i, s] = self.rng.dirichlet(self.gibbs_g[i, s] + np.nansum(f[:, j, s], axis=0))
            st.mark('p_global', i)
        return st, Q_GIBBS, Q_BACK_GIBBS

    def gibbs_sample_p_zones(self, st):
        z = int(self.rng.integers(0, self.Z))
        m = st.source[:, :, 1] & st.zones[z, :, np.newaxis]
        f = m[..., np.newaxis] * self.features
        for i in range(self.F):
            s = self.states[i]
            st.p_zones[z, i, s] = self.rng.dirichlet(1 + np.nansum(f[:, i, s], axis=0))
            st.mark('p_zones', (z, i))
        return st, Q_GIBBS, Q_BACK_GIBBS

    def gibbs_sample_p_families(self, st, frac=0.4):
        k = int(self.rng.integers(0, self.K))
        subset = self.rng.random(self.F) < frac
        feats = np.flatnonzero(subset)
        m = st.source[:, subset, 2] & self.families[k, :, np.newaxis]
        f = m[..., np.newaxis] * self.features[:, subset, :]
        for j, i in enumerate(feats):
            s = self.states[i]
            st.p_families[k, i, s] = self.rng.dirichlet(self.gibbs_f[k, i, s] + np.nansum(f[:, j, s], axis=0))
            st.mark('p_families', (k, i))
        return st, Q_GIBBS, Q_BACK_GIBBS

    # -- alter_* (zone_sampling.py:408-612) --------------------------------------------------------------
    def _dirichlet_proposal(self, v, precision):
        alpha = 1 + precision * v
        v_new = self.rng.dirichlet(alpha)
        lq, lqb = orc.dirichlet_proposal_log_q(v, v_new, precision)
        return v_new, lq, lqb

    def alter_weights(self, st):
        new = st.copy()
        i = int(self.rng.integers(0, self.F))
        if self.inheritance:
            pair = list(self.rng.choice(3, size=2, replace=False))
            cur = st.weights[i, pair]
            v, lq, lqb = self._dirichlet_proposal(cur / cur.sum(), self.precision['weights'])
            new.weights[i, pair] = v * cur.sum()
        else:
            v, lq, lqb = self._dirichlet_proposal(st.weights[i, :], self.precision['weights'])
            new.weights[i, :] = v
        st.mark('weights')
        new.mark('weights')
        return new, lq, lqb

    def _alter_row(self, st, table_name, lead, precision, mark_idx):
        new = st.copy()
        i = mark_idx[-1] if isinstance(mark_idx, tuple) else mark_idx
        pair = list(self.rng.choice(np.flatnonzero(self.states[i]), size=2, replace=False))
        row = getattr(st, table_name)[lead][i]
        cur = row[pair]
        v, lq, lqb = self._dirichlet_proposal(cur / cur.sum(), precision)
        getattr(new, table_name)[lead][i][pair] = v * cur.sum()
        st.mark(table_name, mark_idx)
        new.mark(table_name, mark_idx)
        return new, lq, lqb

    def alter_p_global(self, st):
        i = int(self.rng.integers(0, self.F))
        return self._alter_row(st, 'p_global', 0, self.precision['universal'], i)

    def alter_p_zones(self, st):
        z = int(self.rng.integers(0, self.Z))
        i = int(self.rng.integers(0, self.F))
        return self._alter_row(st, 'p_zones', z, self.precision['contact'], (z, i))

    def alter_p_families(self, st):
        k = int(self.rng.integers(0, self.K))
        i = int(self.rng.integers(0, self.F))
        return self._alter_row(st, 'p_families', k, self.precision['inheritance'], (k, i))

    # -- zone moves (zone_sampling.py:704-933) ------------------------------------------------------------
    def _grow_q(self, site, nbrs, n_free):
        q = (1 - self.rho) / n_free
        if nbrs[site]:
            q += self.rho / np.count_nonzero(nbrs)
        return q

    def _finish_zone_move(self, st, new, z, q, q_back, lqb_s):
        st.mark('zones', z)
        new.mark('zones', z)
        log_q, log_q_back = math.log(q), math.log(q_back)
        if self.sample_source:
            new, lq_s, _ = self.gibbs_sample_sources(new, as_gibbs=self.warmup)    # warm-up: :1293-1296
            if self.warmup:
                return new, Q_GIBBS, Q_BACK_GIBBS
            log_q += lq_s
            log_q_back += lqb_s
        return new, log_q, log_q_back

    def grow_zone(self, st):
        new = st.copy()
        occupied = np.any(st.zones, axis=0)
        lqb_s = self._log_q_back_sources(st) if self.sample_source else 0.
        z = int(self.rng.integers(0, self.Z))
        zone = st.zones[z]
        size = np.count_nonzero(zone)
        if size >= self.max_size:
            return st, Q_REJECT, Q_BACK_REJECT
        nbrs = orc.neighbours(zone, occupied, self.adj)
        cand = nbrs if self.rng.random() < self.rho else ~occupied
        if not np.any(cand):
            return st, Q_REJECT, Q_BACK_REJECT
        site = int(self.rng.choice(cand.nonzero()[0]))
        new.zones[z, site] = True
        q = self._grow_q(site, nbrs, np.count_nonzero(~occupied))
        return self._finish_zone_move(st, new, z, q, 1 / (size + 1), lqb_s)

    def shrink_zone(self, st):
        new = st.copy()
        lqb_s = self._log_q_back_sources(st) if self.sample_source else 0.
        z = int(self.rng.integers(0, self.Z))
        zone = st.zones[z]
        size = np.count_nonzero(zone)
        if size <= self.min_size:
            return st, Q_REJECT, Q_BACK_REJECT
        site = int(self.rng.choice(zone.nonzero()[0]))
        new.zones[z, site] = False
        occ_new = np.any(new.zones, axis=0)
        nbrs_back = orc.neighbours(new.zones[z], occ_new, self.adj)
        q_back = self._grow_q(site, nbrs_back, np.count_nonzero(~occ_new))
        if self.warmup:
            q_back = 1 / (size + 1)                                               # :1552, quirk Q3
        return self._finish_zone_move(st, new, z, 1 / size, q_back, lqb_s)

    def swap_zone(self, st):
        new = st.copy()
        occupied = np.any(st.zones, axis=0)
        lqb_s = self._log_q_back_sources(st) if self.sample_source else 0.
        z = int(self.rng.integers(0, self.Z))
        zone = st.zones[z]
        nbrs = orc.neighbours(zone, occupied, self.adj)
        cand = nbrs if self.rng.random() < self.rho else ~occupied
        if not np.any(cand):
            return st, Q_REJECT, Q_BACK_REJECT
        site_new = int(self.rng.choice(cand.nonzero()[0]))
        new.zones[z, site_new] = True
        site_removed = int(self.rng.choice(zone.nonzero()[0]))
        new.zones[z, site_removed] = False
        n_free = np.count_nonzero(~occupied)
        q = self._grow_q(site_new, nbrs, n_free)
        q_back = self._grow_q(site_removed, orc.neighbours(zone, occupied, self.adj), n_free)
        return self._finish_zone_move(st, new, z, q, q_back, lqb_s)


def default_operators(inheritance, sample_source, steps=None):
    """MCMC.steps_per_operator (mcmc_setup.py:70-95) for the default STEPS of default_config.json:13-20
    after verify_config's normalisation (experiment_setup.py:223-237)."""
    steps = dict(steps or {'area': 0.05, 'weights': 0.4, 'universal': 0.05, 'contact': 0.4, 'inheritance': 0.1,
                           'source': 0.0})
    if not inheritance:
        steps['inheritance'] = 0.0
    if not sample_source:
        steps['source'] = 0.0
    tot = sum(steps.values())
    steps = {k: v / tot for k, v in steps.items()}
    ops = {'shrink_zone': steps['area'] * 0.4, 'grow_zone': steps['area'] * 0.4, 'swap_zone': steps['area'] * 0.2}
    if sample_source:
        ops.update({'gibbs_sample_sources': steps['source'], 'gibbs_sample_weights': steps['weights'],
                    'gibbs_sample_p_global': steps['universal'], 'gibbs_sample_p_zones': steps['contact']})
        if inheritance:
            ops['gibbs_sample_p_families'] = steps['inheritance']
    else:
        ops.update({'alter_weights': steps['weights'], 'alter_p_global': steps['universal'],
                    'alter_p_zones': steps['contact']})
        if inheritance:
            ops['alter_p_families'] = steps['inheritance']
    return ops
