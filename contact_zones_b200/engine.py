"""`Engine`: one C-ABI context (one GPU, a shard of chains) with numpy-friendly methods.

This is the only module that talks to libsbayes_b200.so; the reference-shaped objects in `model.py` and
`sampling/` are thin layers over it.  Array conventions at this level are the reference's `Sample` ones
(`zone_sampling.py:49-75`): zones bool (Z,N), weights (F,C), p_global (1,F,S), p_zones (Z,F,S),
p_families (Fam,F,S), source bool (N,F,C) one-hot (or the packed uint8 (N,F) form used on the device).
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import SbDims, SbOperatorCfg, SbPriorCfg, SbError, as_c, ptr


def state_index(features):
    """One-hot bool (N,F,S) -> uint8 (N,F) state index, 255 = NA (all-zero one-hot, `model.py:112`)."""
    features = np.asarray(features)
    if features.ndim == 2:
        return np.ascontiguousarray(features, dtype=np.uint8)
    idx = np.argmax(features, axis=-1).astype(np.uint8)
    idx[features.sum(axis=-1) == 0] = 255
    return idx


def family_index(families, n_sites):
    """bool (Fam,N) -> int16 (N) family id, -1 = none.  A site listed in several (nested) families belongs to the
    LAST of them, as in the reference: `compute_family_likelihood` writes `lh_families[families[fam]]` family by
    family, so the later one overwrites (`model.py:427-431`)."""
    fid = np.full(n_sites, -1, dtype=np.int16)
    if families is None:
        return fid
    fam = np.asarray(families, dtype=bool)
    for k in range(fam.shape[0]):
        fid[fam[k]] = k
    return fid


def pack_source(source, n_components):
    if source is None:
        return None
    source = np.asarray(source)
    if source.ndim == 3:
        return np.argmax(source, axis=-1).astype(np.uint8)
    return np.ascontiguousarray(source, dtype=np.uint8)


class Engine:
    def __init__(self, features, states, families=None, adj_mat=None, cost_matrix=None, *, n_zones, n_chains=1,
                 inheritance=False, sample_source=False, warmup=False, device=0, chain_offset=0, locations=None):
        self.lib = _lib.load()
        self._ctx = C.c_void_p()
        x = state_index(features)
        self.N, self.F = x.shape
        states = np.asarray(states, dtype=bool)
        self.S = states.shape[1]
        self.Z = int(n_zones)
        self.inheritance = bool(inheritance)
        self.sample_source = bool(sample_source)
        self.warmup = bool(warmup)
        self.Fam = int(np.asarray(families).shape[0]) if (inheritance and families is not None) else 0
        self.C = 3 if inheritance else 2
        self.n_chains = int(n_chains)
        self.chain_offset = int(chain_offset)
        self.device = int(device)
        flags = ((_lib.SB_F_INHERITANCE if inheritance else 0) | (_lib.SB_F_SAMPLE_SOURCE if sample_source else 0) |
                 (_lib.SB_F_WARMUP if warmup else 0))
        dims = SbDims(self.N, self.F, self.S, self.Z, self.Fam, self.n_chains, self.chain_offset, flags)
        rc = self.lib.sb_create(C.byref(self._ctx), int(device), C.byref(dims))
        if rc != 0:
            msg = self.lib.sb_last_error(self._ctx).decode() if self._ctx else 'sb_create failed (no CUDA device?)'
            if self._ctx:
                self.lib.sb_destroy(self._ctx)
                self._ctx = C.c_void_p()
            raise SbError(rc, msg)
        fid = family_index(families, self.N) if inheritance else None
        indptr = indices = None
        if adj_mat is not None:
            if isinstance(adj_mat, tuple):
                indptr, indices = adj_mat
            else:
                csr = adj_mat.tocsr()
                csr.sort_indices()
                indptr, indices = csr.indptr, csr.indices
            indptr, indices = as_c(indptr, np.int32), as_c(indices, np.int32)
        cost = as_c(cost_matrix, np.float64)
        xs, st = as_c(x, np.uint8), as_c(states, np.uint8)
        self._check(self.lib.sb_upload_data(self._ctx, ptr(xs), ptr(st), ptr(fid), ptr(indptr), ptr(indices), ptr(cost)))
        if cost is None and locations is not None:
            # implicit Euclidean cost matrix of the geo prior: N x 2 coordinates instead of N x N doubles
            loc = as_c(np.asarray(locations, dtype=np.float64).reshape(self.N, 2), np.float64)
            self._check(self.lib.sb_upload_locations(self._ctx, ptr(loc)))
        self.has_cost = cost is not None or locations is not None
        self.set_priors()

    # -- plumbing ---------------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise SbError(rc, self.lib.sb_last_error(self._ctx).decode())

    def close(self):
        if getattr(self, '_ctx', None):
            self.lib.sb_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        self._check(self.lib.sb_set_stream(self._ctx, C.c_void_p(cuda_stream or 0)))

    def sync(self):
        self._check(self.lib.sb_sync(self._ctx))

    @property
    def kernel_launches(self):
        v = C.c_int64()
        self._check(self.lib.sb_kernel_launches(self._ctx, C.byref(v)))
        return v.value

    # -- priors -----------------------------------------------------------------------------------
    def set_priors(self, size_type='none', geo_type='uniform', geo_scale=1.0, p_global_type='uniform',
                   p_families_type='uniform', gibbs_counts_global=None, alpha_global=None,
                   gibbs_counts_fam=None, alpha_fam=None):
        cfg = SbPriorCfg(_lib.SIZE_TYPES[size_type], _lib.GEO_TYPES[geo_type], float(geo_scale),
                         _lib.DIR_TYPES[p_global_type], _lib.DIR_TYPES[p_families_type])
        a = [as_c(v, np.float64) for v in (gibbs_counts_global, alpha_global, gibbs_counts_fam, alpha_fam)]
        self._check(self.lib.sb_set_priors(self._ctx, C.byref(cfg), *[ptr(v) for v in a]))

    # -- state ------------------------------------------------------------------------------------
    def set_state(self, chain, zones, weights, p_global, p_zones, p_families=None, source=None):
        zones = as_c(np.asarray(zones).reshape(self.Z, self.N), np.uint8)
        w = as_c(np.asarray(weights).reshape(self.F, self.C), np.float64)
        pg = as_c(np.asarray(p_global).reshape(self.F, self.S), np.float64)
        pz = as_c(np.asarray(p_zones).reshape(self.Z, self.F, self.S), np.float64)
        pf = as_c(np.asarray(p_families).reshape(self.Fam, self.F, self.S), np.float64) if self.inheritance else None
        src = pack_source(source, self.C) if self.sample_source else None
        if self.sample_source and src is None:
            raise ValueError('a SAMPLE_SOURCE engine needs a source array')
        self._check(self.lib.sb_set_state(self._ctx, int(chain), ptr(zones), ptr(w), ptr(pg), ptr(pz), ptr(pf), ptr(src)))

    def set_states(self, chain_begin, zones, weights, p_global, p_zones, p_families=None, source=None):
        """Bulk form of `set_state`: arrays stacked along a leading chain axis, one call for all of them."""
        zones = np.asarray(zones)
        n = zones.shape[0]
        zones = as_c(zones.reshape(n, self.Z, self.N), np.uint8)
        w = as_c(np.asarray(weights).reshape(n, self.F, self.C), np.float64)
        pg = as_c(np.asarray(p_global).reshape(n, self.F, self.S), np.float64)
        pz = as_c(np.asarray(p_zones).reshape(n, self.Z, self.F, self.S), np.float64)
        pf = as_c(np.asarray(p_families).reshape(n, self.Fam, self.F, self.S), np.float64) if self.inheritance else None
        src = None
        if self.sample_source:
            if source is None:
                raise ValueError('a SAMPLE_SOURCE engine needs a source array')
            source = np.asarray(source)
            src = as_c(np.argmax(source, axis=-1) if source.ndim == 4 else source, np.uint8).reshape(n, self.N, self.F)
        self._check(self.lib.sb_set_states(self._ctx, int(chain_begin), int(n), ptr(zones), ptr(w), ptr(pg), ptr(pz),
                                           ptr(pf), ptr(src)))

    def get_state(self, chain, one_hot_source=True, fields=None):
        """Device image of chain -> the `Sample` arrays.  `fields` limits the read-back (e.g. ('zones', 'weights'))."""
        want = (lambda k: fields is None or k in fields)
        zones = np.zeros((self.Z, self.N), np.uint8) if want('zones') else None
        w = np.zeros((self.F, self.C)) if want('weights') else None
        pg = np.zeros((1, self.F, self.S)) if want('p_global') else None
        pz = np.zeros((self.Z, self.F, self.S)) if want('p_zones') else None
        pf = np.zeros((self.Fam, self.F, self.S)) if (self.inheritance and want('p_families')) else None
        src = np.zeros((self.N, self.F), np.uint8) if (self.sample_source and want('source')) else None
        self._check(self.lib.sb_get_state(self._ctx, int(chain), ptr(zones), ptr(w), ptr(pg), ptr(pz), ptr(pf), ptr(src)))
        if src is not None and one_hot_source:
            src = np.eye(self.C, dtype=bool)[src]
        return {'zones': None if zones is None else zones.astype(bool), 'weights': w, 'p_global': pg, 'p_zones': pz,
                'p_families': pf, 'source': src}

    def copy_state(self, dst, src):
        self._check(self.lib.sb_copy_state(self._ctx, int(dst), int(src)))

    # -- evaluation -------------------------------------------------------------------------------
    def eval(self, chain_begin=0, n_chains=None, use_source=None):
        n = self.n_chains - chain_begin if n_chains is None else int(n_chains)
        use_source = self.sample_source if use_source is None else bool(use_source)
        ll, pr, geo = np.zeros(n), np.zeros(n), np.zeros(n)
        self._check(self.lib.sb_eval(self._ctx, int(chain_begin), n, int(use_source), ptr(ll), ptr(pr), ptr(geo)))
        return ll, pr, geo

    def eval_likelihood_async(self, chain_begin=0, n_chains=None, use_source=False):
        n = self.n_chains - chain_begin if n_chains is None else int(n_chains)
        self._check(self.lib.sb_eval_likelihood_async(self._ctx, int(chain_begin), n, int(bool(use_source))))

    def eval_likelihood_noalloc(self, chain_begin=0, n_chains=None, device_ptr=None):
        """`sb_eval_likelihood_noalloc`: the source-free likelihood with the kernel's scratch in the library's static
        device array (no cudaMalloc in the call).  Returns the host array, or writes `n_chains` doubles to the
        DEVICE address `device_ptr` (stream-ordered, no synchronisation) and returns None."""
        n = self.n_chains - chain_begin if n_chains is None else int(n_chains)
        if device_ptr is not None:
            self._check(self.lib.sb_eval_likelihood_noalloc(self._ctx, int(chain_begin), n, C.c_void_p(int(device_ptr))))
            return None
        ll = np.zeros(n)
        self._check(self.lib.sb_eval_likelihood_noalloc(self._ctx, int(chain_begin), n, ptr(ll)))
        return ll

    def read_ll(self, chain_begin=0, n_chains=None):
        n = self.n_chains - chain_begin if n_chains is None else int(n_chains)
        ll = np.zeros(n)
        self._check(self.lib.sb_read_ll(self._ctx, int(chain_begin), n, ptr(ll)))
        return ll

    def components(self, chain):
        lh = np.zeros((self.N, self.F, self.C))
        wn = np.zeros((self.N, self.F, self.C))
        self._check(self.lib.sb_eval_components(self._ctx, int(chain), ptr(lh), ptr(wn)))
        return lh, wn

    def source_posterior(self, chain):
        post = np.zeros((self.N, self.F, self.C))
        self._check(self.lib.sb_source_posterior(self._ctx, int(chain), ptr(post)))
        return post

    def counts(self, chain, recount=False):
        cg = np.zeros((self.F, self.S), np.int32)
        cz = np.zeros((self.Z, self.F, self.S), np.int32)
        cf = np.zeros((self.Fam, self.F, self.S), np.int32) if self.inheritance else None
        csz = np.zeros((self.F, self.C), np.int32)
        csf = np.zeros((self.F, self.C), np.int32)
        self._check(self.lib.sb_counts(self._ctx, int(chain), int(recount), ptr(cg), ptr(cz), ptr(cf), ptr(csz), ptr(csf)))
        return {'cnt_glob': cg, 'cnt_zone': cz, 'cnt_fam': cf, 'cnt_src_zone': csz, 'cnt_src_fam': csf}

    def recount_async(self, chain_begin=0, n_chains=None):
        n = self.n_chains - chain_begin if n_chains is None else int(n_chains)
        self._check(self.lib.sb_recount_async(self._ctx, int(chain_begin), n))

    def device_scalars(self):
        """(ll, prior) of the resident chains as zero-copy torch views of the context's device arrays."""
        import torch
        a, b = C.c_void_p(), C.c_void_p()
        self._check(self.lib.sb_device_scalars(self._ctx, C.byref(a), C.byref(b)))

        class _Dev:
            def __init__(self, addr, n):
                self.__cuda_array_interface__ = {'shape': (n,), 'typestr': '<f8', 'data': (addr, False), 'version': 2}
        dev = torch.device('cuda', self.device)
        return (torch.as_tensor(_Dev(a.value, self.n_chains), device=dev),
                torch.as_tensor(_Dev(b.value, self.n_chains), device=dev))

    def move_log_q(self, chain, kind, zone, site_new=-1, site_removed=-1, p_grow_connected=0.85):
        lq, lqb = C.c_double(), C.c_double()
        self._check(self.lib.sb_move_log_q(self._ctx, int(chain), _lib.OP_IDS[kind + '_zone'], int(zone), int(site_new),
                                           int(site_removed), float(p_grow_connected), C.byref(lq), C.byref(lqb)))
        return lq.value, lqb.value

    def pair_log_q(self, c0, c1, n0, n1, precision):
        """log_q - log_q_back of `dirichlet_proposal` (zone_sampling.py:538-569) for a forced new pair, evaluated by the
        step kernel's own pair-proposal code."""
        out = C.c_double()
        self._check(self.lib.sb_pair_log_q(self._ctx, float(c0), float(c1), float(n0), float(n1), float(precision),
                                           C.byref(out)))
        return out.value

    def rng_sample(self, kind, n, alpha=1.0, seed=0):
        """n draws of the device generator: kind 'uniform', 'gamma', 'normal' (the Gamma sampler's candidates) or
        'flog' (K6 parity surface)."""
        out = np.zeros(int(n))
        self._check(self.lib.sb_rng_sample(self._ctx, {'uniform': 0, 'gamma': 1, 'flog': 2, 'normal': 3}[kind], float(alpha), int(seed), int(n),
                                           ptr(out)))
        return out

    # -- sampler ----------------------------------------------------------------------------------
    def set_operators(self, operators, precision, min_size, max_size, p_grow_connected, seed=0,
                      gibbs_feature_fraction=0.4):
        """operators: {name: weight} as built by `MCMC.steps_per_operator` (`mcmc_setup.py:70-95`);
        precision: {'weights','universal','contact','inheritance'} (`PROPOSAL_PRECISION`);
        max_size / p_grow_connected: scalar or per-chain sequence."""
        cfg = SbOperatorCfg()
        items = [(k, float(v)) for k, v in operators.items() if k in _lib.OP_IDS]
        unknown = [k for k, v in operators.items() if k not in _lib.OP_IDS and float(v) > 0]
        if unknown:
            raise ValueError(f'operators not implemented on the device: {unknown}')
        cfg.n_ops = len(items)
        for i, (k, v) in enumerate(items):
            cfg.op_ids[i] = _lib.OP_IDS[k]
            cfg.op_prob[i] = v
        cfg.precision_weights = float(precision.get('weights', 15))
        cfg.precision_p_global = float(precision.get('universal', 40))
        cfg.precision_p_zones = float(precision.get('contact', 20))
        cfg.precision_p_families = float(precision.get('inheritance') or 20)
        cfg.min_size = int(min_size)
        cfg.gibbs_feature_fraction = float(gibbs_feature_fraction)
        cfg.seed = int(seed) & (2 ** 64 - 1)
        pg = np.ascontiguousarray(np.broadcast_to(np.asarray(p_grow_connected, dtype=np.float64), (self.n_chains,)))
        ms = np.ascontiguousarray(np.broadcast_to(np.asarray(max_size, dtype=np.int32), (self.n_chains,)))
        self._check(self.lib.sb_set_operators(self._ctx, C.byref(cfg), ptr(pg), ptr(ms)))
        self.op_names = [k for k, _ in items]

    def run(self, n_steps):
        self._check(self.lib.sb_run(self._ctx, int(n_steps)))

    def set_sample_from_prior(self, on=True):
        self._check(self.lib.sb_set_sample_from_prior(self._ctx, int(bool(on))))

    def run_events(self, n_steps, event_steps, event_slots, ring_chain=-1):
        """One persistent launch of `n_steps` steps with host-visible events inside it: after `event_steps[k]` completed
        steps every chain records (ll, prior); `event_slots[k]` >= 0 also snapshots the state of `ring_chain` (-1 = of
        every chain) into that slot of the device ring.  Asynchronous."""
        es = np.ascontiguousarray(event_steps, dtype=np.int32)
        sl = np.ascontiguousarray(event_slots, dtype=np.int32)
        assert es.shape == sl.shape
        self._check(self.lib.sb_run_events(self._ctx, int(n_steps), int(es.size), ptr(es), ptr(sl), int(ring_chain)))

    def ring_record_bytes(self):
        """Bytes of one snapshot record (sb_step.cuh:snapshot_state): zone ids + weights + the three p tables."""
        a16 = (lambda v: (v + 15) & ~15)
        fs = self.F * self.S
        return (a16((self.N + 15) & ~15) + a16(self.F * 3 * 8) + a16(fs * 8) + a16(self.Z * fs * 8) +
                a16(self.Fam * fs * 8))

    def event_scalars(self, n_events):
        ll, pr = np.zeros((n_events, self.n_chains)), np.zeros((n_events, self.n_chains))
        self._check(self.lib.sb_read_event_scalars(self._ctx, int(n_events), ptr(ll), ptr(pr)))
        return ll, pr

    def read_snapshot(self, slot, chain):
        zones = np.zeros((self.Z, self.N), np.uint8)
        w = np.zeros((self.F, self.C))
        pg = np.zeros((1, self.F, self.S))
        pz = np.zeros((self.Z, self.F, self.S))
        pf = np.zeros((self.Fam, self.F, self.S)) if self.inheritance else None
        self._check(self.lib.sb_read_snapshot(self._ctx, int(slot), int(chain), ptr(zones), ptr(w), ptr(pg), ptr(pz), ptr(pf)))
        return {'zones': zones.astype(bool), 'weights': w, 'p_global': pg, 'p_zones': pz, 'p_families': pf,
                'source': None}

    def snapshot_ring(self):
        """The device snapshot ring as a zero-copy torch view (n_records, record_bytes) of uint8 -- record (slot, chain)
        is row slot * n_chains + chain -- or None before the first `run_events` with a snapshot slot."""
        import torch
        a, rec, n = C.c_void_p(), C.c_int64(), C.c_int64()
        off, ln = np.zeros(5, np.int64), np.zeros(5, np.int64)
        self._check(self.lib.sb_snapshot_ring(self._ctx, C.byref(a), C.byref(rec), C.byref(n), ptr(off), ptr(ln)))
        self._ring_layout, self.ring_record_size = (off, ln), int(rec.value)
        if not a.value or n.value == 0:
            return None

        class _Dev:
            def __init__(self, addr, shape):
                self.__cuda_array_interface__ = {'shape': shape, 'typestr': '|u1', 'data': (addr, False), 'version': 2}
        return torch.as_tensor(_Dev(a.value, (int(n.value), int(rec.value))), device=torch.device('cuda', self.device))

    def unpack_record(self, rec):
        """One snapshot record (uint8 array, `snapshot_ring` row) -> the arrays `read_snapshot` returns."""
        off, ln = self._ring_layout
        zid = rec[off[0]:off[0] + self.N].view(np.int8)
        zones = zid[None, :] == np.arange(self.Z, dtype=np.int8)[:, None]

        def f64(a, shape):
            return rec[off[a]:off[a] + int(np.prod(shape)) * 8].view(np.float64).reshape(shape).copy()
        fs = (self.F, self.S)
        return {'zones': zones, 'weights': f64(1, (self.F, 3))[:, :self.C].copy(), 'p_global': f64(2, (1,) + fs),
                'p_zones': f64(3, (self.Z,) + fs), 'p_families': f64(4, (self.Fam,) + fs) if self.inheritance else None,
                'source': None}

    def chain_errors(self):
        e = np.zeros(self.n_chains, np.int32)
        self._check(self.lib.sb_chain_errors(self._ctx, ptr(e)))
        return e

    def chain_scalars(self):
        ll, pr = np.zeros(self.n_chains), np.zeros(self.n_chains)
        self._check(self.lib.sb_read_chain_scalars(self._ctx, ptr(ll), ptr(pr)))
        return ll, pr

    def op_stats(self):
        acc = np.zeros((self.n_chains, _lib.SB_N_OPS), np.int64)
        rej = np.zeros((self.n_chains, _lib.SB_N_OPS), np.int64)
        self._check(self.lib.sb_read_op_stats(self._ctx, ptr(acc), ptr(rej)))
        return acc, rej
