"""TEST INFRASTRUCTURE ONLY -- CPU (numpy/scipy) restatement of the sBayes MCMC inner loop.

This module is the *oracle* for the CUDA path in `contact_zones_b200/csrc`.  It may be imported only by
`tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of `bench.py`; the
product (`contact_zones_b200`) never imports it and has no CPU fallback.

Every function restates one piece of the reference's algorithm in dense numpy (same arithmetic, same
O(N*F*C) cost structure, so it doubles as the timed CPU baseline where `/root/reference` is absent)
and cites the reference lines it follows.  Citations are relative to `/root/reference/`.

Pinned: validated against the imported reference itself by `oracle/make_golden.py` (run in the build
container, where `/root/reference` exists); the resulting vectors are committed in `tests/golden/` and
re-checked by `tests/test_oracle_golden.py` on every run.  The reference's own known answers
(`test/test_model.py:52-92`, `util.py:1212-1215` log_binom, `util.py:1098-1103` normalize) are checked
in `tests/test_oracle_reference_tests.py`.

Layout conventions (those of the reference):
    features   bool (N, F, S) one-hot, all-zero row = NA      (model.py:111-112)
    zones      bool (Z, N)                                    (zone_sampling.py:53-54)
    families   bool (Fam, N)
    weights    f64  (F, C)   C = 2 [universal, contact] or 3 [+ inheritance]   (model.py:243-245,289)
    p_global   f64  (1, F, S);  p_zones (Z, F, S);  p_families (Fam, F, S)
    source     bool (N, F, C) one-hot or None
    states     bool (F, S) applicable states
"""
import math

import numpy as np
from scipy.sparse.csgraph import minimum_spanning_tree, csgraph_from_dense
from scipy.special import betaln, gammaln, xlogy

NEG_INF = -np.inf


# ----------------------------------------------------------------------------------------------
# data helpers
# ----------------------------------------------------------------------------------------------

def state_index(features):
    """One-hot (N,F,S) -> uint8 state index (N,F), 255 = NA (all-zero one-hot, model.py:112)."""
    features = np.asarray(features)
    idx = np.argmax(features, axis=-1).astype(np.uint8)
    idx[features.sum(axis=-1) == 0] = 255
    return idx


def one_hot(idx, n_states):
    """Inverse of `state_index`."""
    idx = np.asarray(idx)
    out = np.zeros(idx.shape + (n_states,), dtype=bool)
    ok = idx != 255
    out[ok, idx[ok].astype(np.int64)] = True
    return out


def normalize(x, axis=-1):
    """util.py:1087-1105."""
    return x / np.sum(x, axis=axis, keepdims=True)


def log_binom(n, k):
    """util.py:1202-1217:  log C(n,k) = -betaln(1+n-k, 1+k) - log(n+1)."""
    return -betaln(1 + n - k, 1 + k) - np.log(n + 1)


def dirichlet_logpdf(x, alpha):
    """scipy.stats.dirichlet._logpdf as called by util.py:29-30 (model.py:1168,1216; zone_sampling.py:554,557):
    lgamma(sum a) - sum lgamma(a) + sum xlogy(a-1, x)."""
    x = np.asarray(x, dtype=float)
    alpha = np.asarray(alpha, dtype=float)
    lnB = np.sum(gammaln(alpha)) - gammaln(np.sum(alpha))
    return -lnB + np.sum(xlogy(alpha - 1, x))


# ----------------------------------------------------------------------------------------------
# likelihood  (model.py:145-184, 230-294, 297-452)
# ----------------------------------------------------------------------------------------------

def has_components(zones, families, inheritance):
    """model.py:251-281: (N, C) bool [1, in any zone(, in any family)]."""
    n_sites = zones.shape[1]
    cols = [np.ones(n_sites, dtype=bool), np.any(zones, axis=0)]
    if inheritance:
        cols.append(np.any(np.asarray(families, dtype=bool), axis=0))
    return np.array(cols).T


def normalized_weights(weights, has):
    """model.py:436-452 with the broadcast of :291: (N, F, C)."""
    w = weights[np.newaxis, :, :] * has[:, np.newaxis, :]
    return w / w.sum(axis=2, keepdims=True)


def component_likelihoods(features, zones, families, p_global, p_zones, p_families, inheritance):
    """model.py:230-249 (full rebuild, caching=False): (N, F, C); NA cells = 1 in every component (:247).

    global  :297-330   lh[n,f] = features[n,f,:] . p_global[0,f,:]
    zone    :333-392   lh[zones[z], :] = einsum(features[zones[z]], p_zones[z]); later zones overwrite
    family  :395-433   lh[families[fam], f] = features[families[fam], f, :] . p_families[fam, f, :]
    Rows of sites in no zone / family stay 0 (they get weight 0).
    """
    features = np.asarray(features)
    n_sites, n_features, _ = features.shape
    lh_global = np.einsum('nfs,fs->nf', features, p_global[0], optimize=False).astype(float)
    lh_zone = np.zeros((n_sites, n_features))
    for z in range(len(zones)):
        lh_zone[zones[z], :] = np.einsum('ijk,jk->ij', features[zones[z]], p_zones[z])
    comps = [lh_global, lh_zone]
    if inheritance:
        fam = np.asarray(families, dtype=bool)
        lh_fam = np.zeros((n_sites, n_features))
        for k in range(len(fam)):
            lh_fam[fam[k], :] = np.einsum('ijk,jk->ij', features[fam[k]], p_families[k])
        comps.append(lh_fam)
    all_lh = np.array(comps).transpose((1, 2, 0))
    all_lh[features.sum(axis=-1) == 0] = 1.
    return all_lh


def combine(all_lh, weights, source):
    """model.py:173-184."""
    if source is None:
        return float(np.sum(np.log(np.sum(weights * all_lh, axis=2))))
    sel = np.where(np.asarray(source).ravel())
    w_obs = weights.ravel()[sel]
    if np.any(w_obs == 0):
        return NEG_INF
    return float(np.sum(np.log(w_obs * all_lh.ravel()[sel])))


def log_likelihood(features, families, inheritance, zones, weights, p_global, p_zones, p_families,
                   source=None):
    """Likelihood.__call__(sample, caching=False), model.py:145-171."""
    with np.errstate(divide='ignore', invalid='ignore'):
        all_lh = component_likelihoods(features, zones, families, p_global, p_zones, p_families, inheritance)
        w = normalized_weights(weights, has_components(zones, families, inheritance))
        return combine(all_lh, w, source)


def source_posterior(features, families, inheritance, zones, weights, p_global, p_zones, p_families):
    """zone_sampling.py:196-202: normalize(lh_per_component * weights) -> (N, F, C)."""
    all_lh = component_likelihoods(features, zones, families, p_global, p_zones, p_families, inheritance)
    w = normalized_weights(weights, has_components(zones, families, inheritance))
    return normalize(all_lh * w, axis=-1)


# ----------------------------------------------------------------------------------------------
# priors  (model.py:484-505)
# ----------------------------------------------------------------------------------------------

def size_prior(zones, prior_type):
    """ZoneSizePrior.__call__, model.py:932-972."""
    n_sites = zones.shape[1]
    sizes = np.sum(zones, axis=-1)
    if prior_type == 'uniform':
        return float(-np.sum(log_binom(n_sites, sizes)))
    if prior_type == 'quadratic':
        return float(-np.sum(np.log(sizes ** 2)))
    if prior_type == 'none':
        return 0.
    raise ValueError(f'Invalid prior type {prior_type} for size prior')


def geo_prior_cost_based(zones, cost_mat, scale):
    """geo_prior_distance, model.py:1096-1139, including quirk Q1: the value is overwritten per zone
    (:1137) so only the LAST zone counts.  Per zone: MST over the dense cost sub-matrix of the members
    (scipy drops zero-weight MST edges, :1125-1131), expon.logpdf(edge, scale) = -log(scale) - edge/scale,
    mean over the kept edges; no kept edge -> expon.logpdf(0) = -log(scale)."""
    value = None
    for z in zones:
        sub = cost_mat[z][:, z]
        if sub.shape[0] <= 1:
            raise ValueError("Too few locations to compute distance.")
        mst = minimum_spanning_tree(csgraph_from_dense(sub, null_value=np.inf))
        if mst.nnz > 0:
            d = np.asarray(mst.tocsr()[mst.nonzero()]).ravel()
        else:
            d = np.zeros(1)
        value = -math.log(scale) - d / scale
    return float(np.mean(value))


def scale_counts(counts, scale_to):
    """util.py:569-586."""
    eps = np.finfo(float).eps
    s = np.sum(counts, axis=-1)
    s = np.where(s == 0, eps, s)
    factor = scale_to / s
    factor = np.where(factor < 1, factor, 1)
    return counts * factor[..., None]


def dirichlet_prior_tables(counts, scale_to, initial_counts=1.):
    """PGlobalPrior/PFamiliesPrior.parse_attributes for type 'counts' (model.py:580-587, 655-668) with
    counts_to_dirichlet's extra +1 (util.py:609,620).  Returns (gibbs_counts, density_alpha):
    gibbs_counts = initial_counts + scaled counts  (`prior.counts`, used by the Gibbs operators,
    zone_sampling.py:344,394);  density_alpha = gibbs_counts + 1  (the Dirichlet whose pdf is the prior)."""
    c = np.asarray(counts, dtype=float)
    if scale_to is not None:
        c = scale_counts(c, scale_to)
    gibbs = initial_counts + c
    return gibbs, gibbs + 1.


def dirichlet_param_prior(p, alpha, states):
    """prior_p_global_dirichlet / prior_p_families_dirichlet (model.py:1142-1219) summed:
    p, alpha (..., F, S); states bool (F, S).  Sum over leading dims and features of
    Dirichlet.logpdf(p[..., f, states[f]]; alpha[..., f, states[f]])."""
    p = np.asarray(p, dtype=float).reshape((-1,) + p.shape[-2:])
    alpha = np.asarray(alpha, dtype=float).reshape((-1,) + alpha.shape[-2:])
    total = 0.
    for k in range(p.shape[0]):
        for f in range(p.shape[1]):
            idx = states[f]
            total += dirichlet_logpdf(p[k, f, idx], alpha[k, f, idx])
    return float(total)


def log_prior(zones, p_global, p_families, states, inheritance, prior_cfg, cost_mat=None,
              alpha_global=None, alpha_families=None):
    """Prior.__call__, model.py:484-505.  prior_cfg: {'area_size': {'type'}, 'geo': {'type','scale'},
    'weights': {'type'}, 'universal': {'type'}, 'contact': {'type'}, 'inheritance': {'type'}}.
    Weights / contact priors are 0 for the only supported type 'uniform' (:806-842, 868-886)."""
    total = size_prior(zones, prior_cfg['area_size']['type'])
    geo = prior_cfg['geo']['type']
    if geo == 'cost_based':
        total += geo_prior_cost_based(zones, cost_mat, prior_cfg['geo']['scale'])
    elif geo != 'uniform':
        raise ValueError('Geo prior not supported')
    if prior_cfg['weights']['type'] != 'uniform':
        raise ValueError('Invalid prior type for WeightsPrior')
    u = prior_cfg['universal']['type']
    if u == 'counts':
        total += dirichlet_param_prior(p_global, alpha_global[np.newaxis], states)
    elif u != 'uniform':
        raise ValueError('Invalid prior type for PGlobalPrior')
    if prior_cfg['contact']['type'] != 'uniform':
        raise ValueError('Currently only uniform p_zones priors are supported.')
    if inheritance:
        i = prior_cfg['inheritance']['type']
        if i == 'counts':
            total += dirichlet_param_prior(p_families, alpha_families, states)
        elif i != 'uniform':
            raise ValueError('Invalid prior type for PFamiliesPrior')
    return float(total)


# ----------------------------------------------------------------------------------------------
# count tables  (zone_sampling.py:240,257,284,350,372,400)  -- integer, bit-exact parity surface
# ----------------------------------------------------------------------------------------------

def source_counts(source, mask_sites):
    """Sum of source one-hots over the masked sites -> (F, C) ints  (zone_sampling.py:239-240,256-257,283-284).
    NA cells are counted (their source is a normal draw)."""
    return np.sum(np.asarray(source)[mask_sites], axis=0).astype(np.int64)


def state_counts_global(features, source):
    """zone_sampling.py:339-350: cnt[f,s] = sum_n source[n,f,0] * features[n,f,s]."""
    return np.sum(source[:, :, 0, np.newaxis] & features, axis=0).astype(np.int64)


def state_counts_zone(features, source, zone):
    """zone_sampling.py:366-372."""
    m = source[:, :, 1] & zone[:, np.newaxis]
    return np.sum(m[..., np.newaxis] & features, axis=0).astype(np.int64)


def state_counts_family(features, source, family):
    """zone_sampling.py:389-400."""
    m = source[:, :, 2] & family[:, np.newaxis]
    return np.sum(m[..., np.newaxis] & features, axis=0).astype(np.int64)


# ----------------------------------------------------------------------------------------------
# zone-move proposal probabilities  (zone_sampling.py:704-933; util.py:139-155)
# ----------------------------------------------------------------------------------------------

def neighbours(zone, occupied, adj_mat):
    """util.py:139-155."""
    return np.logical_and(adj_mat.dot(zone), ~occupied)


def _grow_q(site_new, zone, occupied, adj_mat, p_connected):
    nb = neighbours(zone, occupied, adj_mat)
    q = (1 - p_connected) / np.count_nonzero(~occupied)
    if nb[site_new]:
        q += p_connected / np.count_nonzero(nb)
    return q


def grow_log_q(zones, z_id, site_new, adj_mat, p_connected):
    """grow_zone :839-857 (without the source terms): (log_q, log_q_back)."""
    occupied = np.any(zones, axis=0)
    q = _grow_q(site_new, zones[z_id], occupied, adj_mat, p_connected)
    q_back = 1 / (np.count_nonzero(zones[z_id]) + 1)
    return math.log(q), math.log(q_back)


def shrink_log_q(zones, z_id, site_removed, adj_mat, p_connected, warmup=False):
    """shrink_zone :903-926; warm-up variant overwrites q_back (:1552, quirk Q3)."""
    size = np.count_nonzero(zones[z_id])
    q = 1 / size
    new = zones.copy()
    new[z_id, site_removed] = False
    q_back = _grow_q(site_removed, new[z_id], np.any(new, axis=0), adj_mat, p_connected)
    if warmup:
        q_back = 1 / (size + 1)
    return math.log(q), math.log(q_back)


def swap_log_q(zones, z_id, site_new, site_removed, adj_mat, p_connected):
    """swap_zone :752-779: q as grow; q_back with the OLD zone's neighbourhood (:753), in which the removed
    site (a member, hence occupied) can never be a neighbour."""
    occupied = np.any(zones, axis=0)
    q = _grow_q(site_new, zones[z_id], occupied, adj_mat, p_connected)
    q_back = _grow_q(site_removed, zones[z_id], occupied, adj_mat, p_connected)
    return math.log(q), math.log(q_back)


def dirichlet_proposal_log_q(v, v_new, precision):
    """dirichlet_proposal, zone_sampling.py:538-569: (log_q, log_q_back) for a given draw."""
    v = np.asarray(v, dtype=float)
    v_new = np.asarray(v_new, dtype=float)
    return (dirichlet_logpdf(v_new, 1 + precision * v), dirichlet_logpdf(v, 1 + precision * v_new))
