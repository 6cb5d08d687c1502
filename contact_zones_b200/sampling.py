"""Drop-in mirror of `sbayes/sampling/zone_sampling.py` + `mcmc_generative.py` for the hot path: `Sample`, `IndexSet`,
`ZoneMCMC`, `ZoneMCMCWarmup` with the reference's constructor / `generate_samples` / `statistics` contract.

What differs underneath: the chains live on the GPU (one C-ABI context per process, `engine.Engine`), `step()` does
not exist on the host -- `generate_samples` launches the persistent kernel K8 (`sb_run`) for as many steps as fit
between two host-visible events (a logged sample, an MC3 swap, the end of the run) and reads back only what the
reference logs.  Randomness on the device is Philox (K6); host-side randomness (initial zones, MC3 swap decisions)
comes from one seeded numpy Generator instead of the reference's two global generators (quirk Q13).
"""
import logging
import math
import os
import time
from collections import defaultdict

import numpy as np

from . import _lib
from . import device as _dev
from .mc3 import ChainShard, swap_chains


class IndexSet(set):
    """A set of dirty indices with an 'everything is dirty' flag (zone_sampling.py:16-46)."""

    def __init__(self, all_i=True):
        super().__init__()
        self.all = all_i

    def clear(self):
        super().clear()
        self.all = False

    def __bool__(self):
        return self.all or len(self) > 0

    def __copy__(self):
        other = IndexSet(all_i=self.all)
        other.update(self)
        return other

    def __deepcopy__(self, memo):
        return self.__copy__()


def _everything():
    return {'zones': IndexSet(), 'weights': True, 'p_global': IndexSet(), 'p_zones': IndexSet(),
            'p_families': IndexSet()}


class Sample(object):
    """Host interchange type of one chain state (zone_sampling.py:49-115): zones bool (Z,N), weights (F,C),
    p_global (1,F,S), p_zones (Z,F,S), p_families (Fam,F,S) or None, source bool (N,F,C) or None."""

    def __init__(self, zones, weights, p_global, p_zones, p_families, source=None, chain=0):
        self.zones = zones
        self.weights = weights
        self.p_global = p_global
        self.p_zones = p_zones
        self.p_families = p_families
        self.source = source
        self.chain = chain
        self.what_changed = {}
        self.everything_changed()

    @classmethod
    def empty_sample(cls):
        return cls(zones=None, weights=None, p_global=None, p_zones=None, p_families=None)

    def everything_changed(self):
        self.what_changed = {'lh': _everything(), 'prior': _everything()}

    def copy(self):
        dup = (lambda a: None if a is None else np.array(a, copy=True))
        new = Sample(dup(self.zones), dup(self.weights), dup(self.p_global), dup(self.p_zones), dup(self.p_families),
                     source=dup(self.source), chain=self.chain)
        new.what_changed = {k: {n: (v.__copy__() if isinstance(v, IndexSet) else v) for n, v in d.items()}
                            for k, d in self.what_changed.items()}
        return new


def get_max_size_list(start, end, n_total, k_groups):
    """util.py:1182-1199: `k_groups` distinct maximum sizes, repeated to cover `n_total` chains."""
    per_group = math.ceil(n_total / k_groups)
    sizes = np.linspace(start=start, stop=end, num=k_groups, endpoint=False, dtype=int)
    return list(np.repeat(sizes, per_group))[:n_total]


def _smoothed_mle(features_subset, applicable):
    """(count + 1 on applicable states) / row sum -- the initial p_* of zone_sampling.py:1063-1170."""
    cnt = np.sum(features_subset, axis=0).astype(float)
    cnt[applicable] += 1
    return cnt / cnt.sum(axis=1, keepdims=True)


class ZoneMCMC(object):
    """`sbayes.sampling.zone_sampling.ZoneMCMC` (zone_sampling.py:118-1268, mcmc_generative.py:15-378)."""

    IS_WARMUP = False
    Q_GIBBS, Q_BACK_GIBBS = -np.inf, 0
    Q_REJECT, Q_BACK_REJECT = 0, -np.inf

    class ZoneError(Exception):
        pass

    def __init__(self, var_proposal, p_grow_connected, initial_size, initial_sample=None, *, model, data, operators,
                 n_chains, mc3=False, swap_period=None, chain_swaps=None, sample_from_prior=False,
                 show_screen_log=False, logger=None, seed=None, **kwargs):
        self.model, self.data = model, data
        self.n_chains = n_chains
        self.chain_idx = list(range(n_chains))
        self.sample_from_prior = sample_from_prior
        self.mc3, self.swap_period, self.chain_swaps = mc3, swap_period, chain_swaps
        self.show_screen_log = show_screen_log
        self.logger = logger if logger is not None else logging.getLogger()
        self.t_start = time.time()
        self.seed = int(np.random.SeedSequence().entropy % (2 ** 63)) if seed is None else int(seed)
        self.rng = np.random.default_rng(self.seed)                       # host decisions, identical on all ranks

        self.features = data.features
        self.applicable_states = np.asarray(data.states, dtype=bool)
        self.n_states_by_feature = np.sum(data.states, axis=-1)
        self.network = data.network
        self.adj_mat = self.network['adj_mat']
        try:
            self.locations = self.network['locations']
        except (KeyError, AttributeError, TypeError):
            self.locations = None
        self.p_grow_connected = p_grow_connected
        self.min_size, self.max_size = model.min_size, model.max_size
        self.initial_size = initial_size
        self.initial_sample = Sample.empty_sample() if initial_sample is None else initial_sample
        self.inheritance = model.inheritance
        self.families = data.families
        self.n_sites, self.n_features = self.features.shape[:2]
        self.n_zones = model.n_zones
        self.n_families = self.families.shape[0] if self.inheritance else None
        self.n_sources = 3 if self.inheritance else 2
        self.var_proposal_weight = var_proposal['weights']
        self.var_proposal_p_global = var_proposal['universal']
        self.var_proposal_p_zones = var_proposal['contact']
        self.var_proposal_p_families = var_proposal.get('inheritance')
        self.var_proposal = dict(var_proposal)

        self.fn_operators, self.p_operators = self.get_operators(operators)
        self.operators = {k: float(v) for k, v in operators.items()}
        self.statistics = {'sample_id': [], 'sample_likelihood': [], 'sample_prior': [], 'sample_zones': [],
                           'sample_weights': [], 'sample_p_global': [], 'sample_p_zones': [], 'sample_p_families': [],
                           'last_sample': [], 'acceptance_ratio': math.nan, 'accepted_steps': 0, 'n_swaps': 0,
                           'accepted_swaps': 0, 'swap_ratio': [], 'accept_operator': defaultdict(int),
                           'reject_operator': defaultdict(int)}
        self._ll = np.full(n_chains, -np.inf)
        self._prior = np.full(n_chains, -np.inf)
        self.shard = ChainShard(n_chains)
        if self.shard.world > 1:            # every rank must take the same host decisions: rank 0's seed wins
            import torch.distributed as dist
            box = [self.seed]
            dist.broadcast_object_list(box, src=0)
            self.seed = int(box[0])
        self._swap_rng = np.random.default_rng([self.seed, 0])
        self._engine = None
        self._dist_device = None

    # -- operators ------------------------------------------------------------------------------------------------
    def get_operators(self, operators):
        """zone_sampling.py:1250-1266: names -> (operators, probabilities).  On the device an operator is an id in the
        step kernel's table, so `fn_operators` holds the names; an unknown name with a non-zero weight is an error
        (the reference fails in `getattr`)."""
        names, probs = [], []
        for k, v in operators.items():
            if k not in _lib.OP_IDS:
                if float(v) != 0.0:
                    raise AttributeError(f"'{type(self).__name__}' object has no attribute '{k}'")
                continue
            names.append(k)
            probs.append(float(v))
        return names, probs

    # -- initial state (host, a18: zone_sampling.py:935-1233) ------------------------------------------------------
    def _neighbours(self, zone, occupied):
        """util.py:139-155: sites adjacent to the zone that belong to no zone."""
        return np.logical_and(self.adj_mat.dot(zone), ~occupied)

    def grow_zone_of_size_k(self, k, already_in_zone=None):
        """zone_sampling.py:992-1031: random connected zone of k free sites; raises ZoneError when stuck."""
        occupied = np.zeros(self.n_sites, bool) if already_in_zone is None else already_in_zone
        free = np.flatnonzero(~occupied)
        if free.size == 0:
            raise self.ZoneError
        zone = np.zeros(self.n_sites, bool)
        seed_site = free[self.rng.integers(free.size)]
        zone[seed_site] = occupied[seed_site] = True
        for _ in range(k - 1):
            cand = np.flatnonzero(self._neighbours(zone, occupied))
            if cand.size == 0:
                raise self.ZoneError
            site = cand[self.rng.integers(cand.size)]
            zone[site] = occupied[site] = True
        return zone, occupied

    def generate_initial_zones(self):
        """zone_sampling.py:935-990: zones of a previous sample are kept, the others are grown (up to 1000 restarts)."""
        zones = np.zeros((self.n_zones, self.n_sites), bool)
        if self.n_zones == 0:
            return zones
        n_given = 0
        occupied0 = np.zeros(self.n_sites, bool)
        if self.initial_sample.zones is not None:
            for i, z in enumerate(self.initial_sample.zones):
                zones[i] = z
                occupied0 |= np.asarray(z, bool)
                n_given += 1
        for _attempt in range(1001):
            occupied = occupied0.copy()
            try:
                for i in range(n_given, self.n_zones):
                    zones[i], occupied = self.grow_zone_of_size_k(self.initial_size, occupied)
                return zones
            except self.ZoneError:
                zones[n_given:] = False
        raise ValueError("Failed to add additional area. Try fewer areas"
                         "or set initial_sample to None")

    def generate_initial_weights(self):
        if self.initial_sample.weights is not None:
            w = np.asarray(self.initial_sample.weights, dtype=float)
        else:
            w = np.full((self.n_features, self.n_sources), 1.)
        return w / w.sum(axis=-1, keepdims=True)

    def generate_initial_p_global(self):
        if self.initial_sample.p_global is not None:
            return np.asarray(self.initial_sample.p_global)
        return _smoothed_mle(self.features, self.applicable_states)[np.newaxis]

    def generate_initial_p_zones(self, initial_zones):
        p = np.zeros((self.n_zones, self.n_features, self.features.shape[2]))
        n_given = 0
        if self.initial_sample.p_zones is not None:
            for i, pz in enumerate(self.initial_sample.p_zones):
                p[i] = pz
                n_given += 1
        for i in range(n_given, self.n_zones):
            p[i] = _smoothed_mle(self.features[initial_zones[i]], self.applicable_states)
        return p

    def generate_initial_p_families(self):
        if self.initial_sample.p_families is not None:
            return np.array(self.initial_sample.p_families, dtype=float)
        fam = np.asarray(self.families, bool)
        return np.stack([_smoothed_mle(self.features[fam[k]], self.applicable_states) for k in range(len(fam))])

    def _initial_sample_host(self, c=0):
        self.rng = np.random.default_rng([self.seed, 1 + int(c)])         # per chain: independent of the sharding
        zones = self.generate_initial_zones()
        return Sample(zones=zones, weights=self.generate_initial_weights(), p_global=self.generate_initial_p_global(),
                      p_zones=self.generate_initial_p_zones(zones),
                      p_families=self.generate_initial_p_families() if self.inheritance else None, chain=c)

    def generate_initial_sample(self, c=0):
        """zone_sampling.py:1197-1233.  With SAMPLE_SOURCE the initial source is one Gibbs sweep -- run on the device
        (operator `gibbs_sample_sources`, K7) through a pooled one-chain context."""
        sample = self._initial_sample_host(c)
        if self.model.sample_source:
            eng = _dev.engine_for(self.data, self.n_zones, self.inheritance, True)
            _dev.upload_sample(eng, 0, sample)
            self._source_sweep(eng, seed=self.seed + 7919 * (c + 1))
            sample.source = eng.get_state(0, fields=('source',))['source']
        return sample

    def _source_sweep(self, eng, seed):
        eng.eval()
        eng.set_operators({'gibbs_sample_sources': 1.0}, self.var_proposal, min_size=self.min_size,
                          max_size=self.max_size if np.isscalar(self.max_size) else max(self.max_size),
                          p_grow_connected=0.5, seed=seed)
        eng.run(1)
        eng.sync()

    # -- device plumbing ------------------------------------------------------------------------------------------
    def _per_chain(self, v):
        """Scalar or per-GLOBAL-chain sequence -> this rank's slice."""
        if np.isscalar(v):
            return v
        return list(v)[self.shard.offset:self.shard.offset + self.shard.n_local]

    def _open_engine(self):
        if self._engine is not None:
            return self._engine
        prior = self.model.prior
        kw = prior.device_config()
        eng = _dev.engine_for(self.data, self.n_zones, self.inheritance, self.model.sample_source,
                              n_chains=self.shard.n_local, with_cost=kw['geo_type'] == 'cost_based',
                              warmup=self.IS_WARMUP, chain_offset=self.shard.offset, pooled=False)
        eng.set_priors(**kw)
        if self.sample_from_prior:          # mcmc_generative.py:115-116: the likelihood is identically 0
            eng.set_sample_from_prior(True)
        self._engine = eng
        if self.shard.world > 1:
            import torch
            import torch.distributed as dist
            self._dist_device = (torch.device('cuda', _dev.default_device())
                                 if dist.get_backend() == 'nccl' else None)
        return eng

    def _set_run_operators(self, eng):
        ops = {k: v for k, v in zip(self.fn_operators, self.p_operators)}
        eng.set_operators(ops, self.var_proposal, min_size=self.min_size, max_size=self._per_chain(self.max_size),
                          p_grow_connected=self._per_chain(self.p_grow_connected), seed=self.seed)

    def _sample_of(self, eng, local_chain, chain, with_source=True):
        st = eng.get_state(local_chain, fields=None if with_source else ('zones', 'weights', 'p_global', 'p_zones',
                                                                           'p_families'))
        return Sample(st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'], source=st['source'],
                      chain=chain)

    def _template(self, with_source):
        z, n, f, s = self.n_zones, self.n_sites, self.n_features, self.features.shape[2]
        t = {'zones': ((z, n), bool), 'weights': ((f, self.n_sources), float), 'p_global': ((1, f, s), float),
             'p_zones': ((z, f, s), float)}
        if self.inheritance:
            t['p_families'] = ((self.n_families, f, s), float)
        if with_source and self.model.sample_source:
            t['source'] = ((n, f, self.n_sources), bool)
        return t

    def _fetch_sample(self, chain, with_source):
        """The host `Sample` of GLOBAL chain `chain` on every rank (its owner broadcasts it)."""
        eng = self._engine
        owner, local = self.shard.owner(chain)

        def local_arrays():
            smp = self._sample_of(eng, local, chain, with_source)
            d = {'zones': smp.zones, 'weights': smp.weights, 'p_global': smp.p_global, 'p_zones': smp.p_zones}
            if self.inheritance:
                d['p_families'] = smp.p_families
            if with_source and self.model.sample_source:
                d['source'] = smp.source
            return d

        a = self.shard.broadcast_state(chain, local_arrays, self._template(with_source), self._dist_device)
        return Sample(a['zones'], a['weights'], a['p_global'], a['p_zones'], a.get('p_families'), source=a.get('source'),
                      chain=chain)

    def _refresh_scalars(self):
        ll, pr = self._engine.chain_scalars()
        if self.sample_from_prior:
            ll = np.zeros_like(ll)
        self._ll, self._prior = self.shard.all_gather_scalars(ll, pr, self._dist_device)

    # events per persistent launch / bytes of snapshot ring per launch (the ring holds one record per logged sample
    # and resident chain under MC3, one per logged sample otherwise)
    MAX_EVENTS_PER_LAUNCH = 4096
    RING_BYTES_PER_LAUNCH = 256 << 20     # (x 2: consecutive launches alternate between two halves of the ring; 1 / 4 GB
                                          # were measured: no faster, and multi-GB allocations stall now and then)

    def _snapshot_samples(self, requests):
        """The logged `Sample`s of a launch: `requests` = [(snapshot slot, GLOBAL chain)], each taken from the ring of the
        chain's owner; every rank gets all of them in one collective on device memory (`ChainShard.gather_records`)."""
        eng = self._engine
        if not requests:
            return []
        ring = eng.snapshot_ring()
        if ring is None:                     # a rank whose chains are never logged (no MC3) has no ring: it only receives
            import torch
            ring = torch.zeros(1, eng.ring_record_size, dtype=torch.uint8, device=torch.device('cuda', eng.device))
        recs = self.shard.gather_records(
            [(chain, slot * self.shard.counts[self.shard.owner(chain)[0]] + self.shard.owner(chain)[1])
             for slot, chain in requests], ring)
        out = []
        for rec, (_, chain) in zip(recs, requests):
            a = eng.unpack_record(rec)
            out.append(Sample(a['zones'], a['weights'], a['p_global'], a['p_zones'], a['p_families'] if self.inheritance else None,
                              source=None, chain=chain))
        return out

    # evaluation API of the reference sampler (mcmc_generative.py:86-126)
    def prior(self, sample, chain):
        return self.model.prior(sample)

    def likelihood(self, sample, chain):
        if self.sample_from_prior:
            return 0.
        return self.model.likelihood(sample)

    # -- the run (a11: mcmc_generative.py:149-240) ----------------------------------------------------------------
    def _initialise_chains(self):
        """Initial samples of this rank's chains -> device; the initial source sweep runs on all chains at once."""
        eng = self._open_engine()
        smp = [self._initial_sample_host(self.shard.offset + lc) for lc in range(self.shard.n_local)]
        if smp:      # one bulk upload (sb_set_states); the sources are redrawn on the device right below
            eng.set_states(0, np.stack([s.zones for s in smp]), np.stack([s.weights for s in smp]),
                           np.stack([s.p_global for s in smp]), np.stack([s.p_zones for s in smp]),
                           np.stack([s.p_families for s in smp]) if self.inheritance else None,
                           np.zeros((len(smp), self.n_sites, self.n_features), np.uint8)
                           if self.model.sample_source else None)
        if self.model.sample_source:          # the reference always redraws the sources (zone_sampling.py:1226-1231)
            self._source_sweep(eng, seed=self.seed ^ 0x5DEECE66D)
        eng.eval()
        self._set_run_operators(eng)
        self._refresh_scalars()
        return eng

    def _collect_operator_stats(self, eng, base):
        acc, rej = eng.op_stats()
        tot = np.stack([acc.sum(axis=0), rej.sum(axis=0)]).astype(np.float64) - base
        if self.shard.world > 1:
            import torch
            import torch.distributed as dist
            t = torch.from_numpy(tot)
            t = t.to(self._dist_device) if self._dist_device is not None else t
            dist.all_reduce(t)
            tot = t.cpu().numpy()
        for name in self.fn_operators:
            i = _lib.OP_IDS[name]
            self.statistics['accept_operator'][name] += int(tot[0, i])
            self.statistics['reject_operator'][name] += int(tot[1, i])
        self.statistics['accepted_steps'] += int(tot[0].sum())

    def generate_samples(self, n_steps, n_samples, warm_up=False, warm_up_steps=None):
        """Run the sampler.  Warm-up mode returns the `Sample` of the chain with the highest posterior
        (mcmc_generative.py:183-200); sampling mode fills `self.statistics` and returns None (:203-240)."""
        eng = self._initialise_chains()
        acc0, rej0 = eng.op_stats()
        base = np.stack([acc0.sum(axis=0), rej0.sum(axis=0)]).astype(np.float64)

        if warm_up:
            print("Tuning parameters in warm-up...")
            done = 0
            for tenth in range(10):
                print("warm-up", tenth * 10, "%")
                upto = (warm_up_steps * (tenth + 1)) // 10
                eng.run(upto - done)
                done = upto
            self._refresh_scalars()
            posterior = self._ll + self._prior
            best_chain = int(np.argmax(posterior))               # list.index(max(...)): first maximum
            return self._fetch_sample(best_chain, with_source=True)

        print("Sampling from posterior...")
        steps_per_sample = int(np.ceil(n_steps / n_samples))
        t_start = time.time()
        # host-visible events, keyed by the number of completed steps t = i_step + 1 (quirk Q12)
        log_at = {i + 1: i // steps_per_sample for i in range(0, n_steps, steps_per_sample)}
        swap_at = set(range(self.swap_period, n_steps + 1, self.swap_period)) if self.mc3 else set()
        screen_at = set(range(1000, n_steps + 1, 1000)) if self.show_screen_log else set()
        stops = sorted(set(log_at) | swap_at | screen_at | {n_steps})
        # The steps between two host-visible events never end the persistent launch: one launch covers as many events
        # as its snapshot ring holds (`sb_run_events`); (ll, prior) of every chain is recorded at each event and the
        # logged chain's state at each logging point.  Without MC3 the logged chain never changes, so only that chain
        # is snapshotted; with MC3 the label moves at the swap points, which are replayed on the host afterwards from
        # the recorded scalars -- every chain snapshots and the host reads the one that carried label 0.
        eng = self._engine
        n_local = max(1, self.shard.n_local)
        per_log = eng.ring_record_bytes() * (n_local if self.mc3 else 1)
        max_logs = max(1, self.RING_BYTES_PER_LAUNCH // max(1, per_log))
        ring_chain = -1
        if not self.mc3:
            owner, local = self.shard.owner(self.chain_idx[0])
            ring_chain = local if owner == self.shard.rank else None
        # the launches of the run: consecutive stops, as many logging points each as the ring holds
        launches, i = [], 0
        while i < len(stops):
            batch, n_logs = [], 0
            while i < len(stops) and len(batch) < self.MAX_EVENTS_PER_LAUNCH:
                t = stops[i]
                if t in log_at and n_logs == max_logs:
                    break
                batch.append(t)
                n_logs += t in log_at
                i += 1
            launches.append(batch)

        def enqueue(b):
            # consecutive launches write alternating halves of the ring, so that launch b + 1 can run while the host
            # reads the snapshots of launch b
            batch, first = launches[b], (launches[b - 1][-1] if b else 0)
            slots, k = [], (b & 1) * max_logs
            for t in batch:
                if t in log_at and ring_chain is not None:
                    slots.append(k)
                    k += 1
                else:
                    slots.append(-1)
            eng.run_events(batch[-1] - first, [t - first for t in batch], slots, -1 if ring_chain is None else ring_chain)

        trace = {} if os.environ.get('SBAYES_B200_TRACE') else None          # seconds per host phase of the run

        def lap(name, t0):
            if trace is not None:
                trace[name] = trace.get(name, 0.0) + time.time() - t0
            return time.time()

        if launches:
            enqueue(0)
        for b, batch in enumerate(launches):
            t0 = time.time()
            ll_ev, pr_ev = eng.event_scalars(len(batch))                   # waits for launch b
            t0 = lap('wait for the launch', t0)
            if b + 1 < len(launches):
                enqueue(b + 1)                                             # ... which runs under the host work below
            t0 = lap('enqueue', t0)
            if self.sample_from_prior:
                ll_ev = np.zeros_like(ll_ev)
            ll_ev, pr_ev = self.shard.all_gather_events(ll_ev, pr_ev, self._dist_device)
            t0 = lap('all-gather of the event scalars', t0)
            # replay of the launch's events in order: the swaps move the labels, so the chain that carries label 0 at
            # every logging point is known only now; its snapshots are then fetched together (one collective per launch)
            k, logged = (b & 1) * max_logs, []
            for j, t in enumerate(batch):
                self._ll, self._prior = ll_ev[j], pr_ev[j]                 # (ll, prior) of every chain after t steps
                if t in log_at:
                    logged.append((j, t, k, self.chain_idx[0]))
                    k += 1
                if t in swap_at:
                    self.swap_chains(None)
                if t in screen_at:
                    self.print_screen_log(t, None)
            t0 = lap('replay of swaps', t0)
            for (j, t, _, c0), sample in zip(logged, self._snapshot_samples([(k_, c0) for _, _, k_, c0 in logged])):
                self._ll, self._prior = ll_ev[j], pr_ev[j]
                self.log_sample_statistics(sample, c=c0, sample_id=log_at[t])
            self._ll, self._prior = ll_ev[-1], pr_ev[-1]
            lap('snapshots', t0)
        if trace is not None:
            print('generate_samples host phases [s]:', {k: round(v, 4) for k, v in trace.items()}, 'launches', len(launches))
        if n_steps > 1:
            self.log_last_sample(self._fetch_sample(self.chain_idx[0], with_source=True))
        eng.sync()
        t_end = time.time()
        self._collect_operator_stats(eng, base)
        st = self.statistics
        st['sampling_time'] = t_end - t_start
        st['time_per_sample'] = (t_end - t_start) / n_samples
        st['acceptance_ratio'] = st['accepted_steps'] / n_steps             # quirk Q4: summed over all chains
        st['acceptance_ratio_per_chain'] = st['accepted_steps'] / (n_steps * self.n_chains)
        st['swap_ratio'] = st['accepted_swaps'] / st['n_swaps'] if st['n_swaps'] > 0 else 0

    def swap_chains(self, sample):
        """mcmc_generative.py:242-280 on the gathered (ll, prior): labels move, states stay where they are."""
        swap_chains(self.chain_idx, self._ll, self._prior, self.chain_swaps, self._swap_rng, self.statistics)

    def log_sample_statistics(self, sample, c, sample_id):
        st = self.statistics
        st['sample_id'].append(sample_id)
        st['sample_zones'].append(sample.zones)
        st['sample_weights'].append(sample.weights)
        st['sample_p_global'].append(sample.p_global)
        st['sample_p_zones'].append(sample.p_zones)
        st['sample_p_families'].append(sample.p_families)
        st['sample_likelihood'].append(float(self._ll[c]))
        st['sample_prior'].append(float(self._prior[c]))
        if self.show_screen_log:
            print('Log-likelihood: %.2f' % self._ll[c])

    def log_last_sample(self, last_sample):
        self.statistics['last_sample'] = last_sample

    def print_screen_log(self, i_step, sample):
        per_million = (time.time() - self.t_start) / (i_step + 1) * 1000000
        print(str(i_step).ljust(12) + ('log-likelihood:  %.2f' % self._ll[self.chain_idx[0]]).ljust(36) +
              '%i seconds / million steps' % per_million)

    def print_statistics(self, samples):
        self.logger.info("\n")
        self.logger.info("MCMC STATISTICS")
        self.logger.info("##########################################")
        self.logger.info(log_operator_statistics_header())
        for name in self.fn_operators:
            self.logger.info(log_operator_statistics(name, samples))

    def close(self):
        if self._engine is not None:
            self._engine.close()
            self._engine = None


class ZoneMCMCWarmup(ZoneMCMC):
    """zone_sampling.py:1272-1577: per-chain `max_size` (4 groups) and `p_grow_connected` in {0.95, cfg}; the device
    runs the warm-up variants of the operators (SB_F_WARMUP: forced-accept source moves :1293-1296, quirk Q3)."""

    IS_WARMUP = True

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.max_size = [int(v) for v in get_max_size_list(start=(self.initial_size + self.max_size) / 4,
                                                           end=self.max_size, n_total=self.n_chains, k_groups=4)]
        pick = np.random.default_rng([self.seed, 2]).integers(0, 2, size=self.n_chains)
        self.p_grow_connected = [0.95 if p == 0 else self.p_grow_connected for p in pick]


COL_WIDTHS = [20, 8, 8, 8, 10]


def log_operator_statistics_header():
    cols = ['OPERATOR', 'ACCEPTS', 'REJECTS', 'TOTAL']
    return '\t'.join([c.ljust(COL_WIDTHS[i]) for i, c in enumerate(cols)] + ['ACC. RATE'])


def log_operator_statistics(operator_name, mcmc_stats):
    acc = mcmc_stats['accept_operator'][operator_name]
    rej = mcmc_stats['reject_operator'][operator_name]
    total = acc + rej
    if total == 0:
        return '\t'.join(x.ljust(COL_WIDTHS[i]) for i, x in enumerate([operator_name, '-', '-', '-', '-']))
    cells = [operator_name, str(acc), str(rej), str(total)]
    return '\t'.join([c.ljust(COL_WIDTHS[i]) for i, c in enumerate(cells)] + ['%.2f%%' % (100 * acc / total)])
