"""TEST INFRASTRUCTURE ONLY -- generate `tests/golden/*.npz` by running the UNMODIFIED reference
(imported from /root/reference through `oracle/refshim.py`) on seeded synthetic inputs, and check the
numpy restatement (`oracle/sbayes_oracle.py`) against it at the same time.

Run in the build container (where /root/reference exists):   python oracle/make_golden.py
Each case stores the inputs (data + state) and the reference's outputs:
  ll_nosrc / ll_src       Likelihood.__call__(sample, caching=False)            model.py:145
  prior, geo, size        Prior.__call__ and its terms                          model.py:484,932,1018
  post                    source posterior normalize(lh * weights)              zone_sampling.py:196-202
  cnt_*                   the integer count tables of the Gibbs operators       zone_sampling.py:240-400
  moves                   (kind, z, site_new, site_removed, log_q, log_q_back) from the reference's
                          grow/shrink/swap operators (resample_source=False)    zone_sampling.py:704-933
"""
import os
import random
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import refshim  # noqa: E402
import sbayes_oracle as orc  # noqa: E402
from contact_zones_b200 import synthetic  # noqa: E402

ref = refshim.modules()

CASES = {
    # name: (N, F, S, Fam, Z, zone_size, inheritance, ragged, counts_prior, geo, size_prior)
    'A_sim_like':    (120, 12, 4, 0, 1, 10, False, False, False, 'uniform', 'none'),
    'B_balkan_like': (28, 47, 3, 5, 3, 4, True, True, True, 'uniform', 'uniform'),
    'C_samerica_like': (100, 36, 5, 6, 3, 6, True, True, True, 'cost_based', 'quadratic'),
    'D_small':       (300, 40, 8, 7, 4, 12, True, False, False, 'cost_based', 'uniform'),
    'E_small':       (500, 20, 4, 0, 4, 25, False, False, False, 'uniform', 'none'),
}
# the reference's BUNDLED data sets, read with the reference's own loaders (oracle/refdata.py):
# name: (experiment, Z, zone_size, geo, size_prior) -- priors / scale_counts as in the experiment's config.json
BUNDLED_CASES = {
    'F_balkan_data':   ('balkan', 3, 4, 'uniform', 'uniform'),
    'G_samerica_data': ('south_america', 5, 6, 'cost_based', 'quadratic'),
}


def prior_config(inheritance, counts_prior, geo, size, scale_u=10., scale_f=10., geo_scale=100.):
    cfg = {'geo': {'type': geo, 'scale': geo_scale}, 'area_size': {'type': size}, 'weights': {'type': 'uniform'},
           'universal': {'type': 'counts', 'scale_counts': scale_u} if counts_prior else {'type': 'uniform'},
           'contact': {'type': 'uniform'}}
    if inheritance:
        cfg['inheritance'] = ({'type': 'counts', 'scale_counts': scale_f} if counts_prior else {'type': 'uniform'})
    return cfg


def rel_err(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(1e-300, np.maximum(np.abs(a), np.abs(b)))))


def run_case(name, spec, seed):
    scale_u = scale_f = 10.
    geo_scale = 100.
    if name in BUNDLED_CASES:
        import refdata
        exp, z, zs, geo, size = spec
        data, info = refdata.load_bundled(exp)
        n, f, s = data.features.shape
        fam, inh, ragged, cp = data.families.shape[0], info['inheritance'], True, True
        scale_u, scale_f = info['scale_universal'], info['scale_inheritance']
        if geo == 'cost_based':        # scale of the order of a Delaunay edge, so that the MST term is O(1)
            geo_scale = float(np.median(data.network['dist_mat'][data.network['adj_mat'].toarray() > 0]))
    else:
        n, f, s, fam, z, zs, inh, ragged, cp, geo, size = spec
        data = synthetic.make_data(n, f, s, n_families=fam, na_frac=0.03, seed=seed, ragged_states=ragged,
                                   counts_prior=cp)
    raw_univ = np.array(data.prior_universal['counts'], float) if cp else None
    raw_fam = np.array(data.prior_inheritance['counts'], float) if (cp and inh) else None
    st = synthetic.random_state(data, z, zs, inh, with_source=True, seed=seed + 100)
    pcfg = prior_config(inh, cp, geo, size, scale_u, scale_f, geo_scale)
    model_cfg = {'N_AREAS': z, 'MIN_M': 3, 'MAX_M': 50, 'INHERITANCE': inh, 'SAMPLE_SOURCE': True, 'PRIOR': pcfg}
    model = ref.model.Model(data, model_cfg)   # NB scales data.prior_*['counts'] in place (quirk Q10)
    Sample = ref.zone_sampling.Sample

    def sample(src):
        return Sample(zones=st['zones'].copy(), weights=st['weights'].copy(), p_global=st['p_global'].copy(),
                      p_zones=st['p_zones'].copy(),
                      p_families=None if st['p_families'] is None else st['p_families'].copy(),
                      source=None if not src else st['source'].copy())

    out = {}
    out['ll_nosrc'] = model.likelihood(sample(False), caching=False)
    out['ll_src'] = model.likelihood(sample(True), caching=False)
    smp = sample(False)
    out['prior'] = model.prior(smp)
    out['geo'] = model.prior.geo_prior.cached
    out['size'] = model.prior.size_prior.cached
    lik = model.likelihood
    s2 = sample(True)
    lh = lik.update_component_likelihoods(sample=s2, caching=False)
    w = lik.update_weights(sample=s2)
    out['post'] = ref.util.normalize(lh * w, axis=-1)
    out['wnorm'] = w.copy()
    has_zone = st['zones'].any(0)
    src = st['source']
    out['cnt_src_zone'] = np.sum(src[has_zone], axis=0)
    if inh:
        out['cnt_src_fam'] = np.sum(src[lik.has_family], axis=0)
        out['cnt_fam'] = np.stack([
            np.nansum((src[:, :, 2] & data.families[k, :, None])[..., None] * data.features, axis=0)
            for k in range(fam)])
    out['cnt_glob'] = np.nansum(src[:, :, 0, None] * data.features, axis=0)
    out['cnt_zone'] = np.stack([
        np.nansum((src[:, :, 1] & st['zones'][k, :, None])[..., None] * data.features, axis=0) for k in range(z)])

    # ---- oracle vs reference ----
    fam_b = np.asarray(data.families, bool)
    args = (data.features, fam_b, inh, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'])
    errs = {
        'll_nosrc': rel_err(orc.log_likelihood(*args, source=None), out['ll_nosrc']),
        'll_src': rel_err(orc.log_likelihood(*args, source=src), out['ll_src']),
        'post': float(np.max(np.abs(orc.source_posterior(*args) - out['post']))),
    }
    ag = af = None
    gibbs_g = gibbs_f = None
    if cp:
        gibbs_g, ag = orc.dirichlet_prior_tables(raw_univ, scale_u)
        assert np.array_equal(gibbs_g, model.prior.prior_p_global.counts)
        if inh:
            gibbs_f, af = orc.dirichlet_prior_tables(raw_fam, scale_f)
            assert np.array_equal(gibbs_f, model.prior.prior_p_families.counts)
    o_prior = orc.log_prior(st['zones'], st['p_global'], st['p_families'], data.states, inh, pcfg,
                            cost_mat=data.geo_prior['cost_matrix'], alpha_global=ag, alpha_families=af)
    errs['prior'] = rel_err(o_prior, out['prior'])
    assert np.array_equal(orc.source_counts(src, has_zone), out['cnt_src_zone'])
    assert np.array_equal(orc.state_counts_global(data.features, src), out['cnt_glob'])
    for k in range(z):
        assert np.array_equal(orc.state_counts_zone(data.features, src, st['zones'][k]), out['cnt_zone'][k])
    if inh:
        for k in range(fam):
            assert np.array_equal(orc.state_counts_family(data.features, src, fam_b[k]), out['cnt_fam'][k])

    # ---- zone-move proposal probabilities from the reference's own operators ----
    ops = {'grow_zone': 1.}
    sampler = ref.zone_sampling.ZoneMCMC(
        data=data, model=model, n_chains=1, operators=ops,
        var_proposal={'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20},
        p_grow_connected=0.85, initial_size=5)
    moves = []
    random.seed(seed)
    np.random.seed(seed)
    for kind in ['grow', 'shrink', 'swap'] * 8:
        smp = sample(True)
        op = getattr(sampler, kind + '_zone')
        new, lq, lqb = op(smp, resample_source=False)
        if new is smp:
            continue
        diff = new.zones.astype(int) - smp.zones.astype(int)
        zz = int(np.flatnonzero(np.abs(diff).sum(1))[0])
        added = np.flatnonzero(diff[zz] == 1)
        removed = np.flatnonzero(diff[zz] == -1)
        sn = int(added[0]) if len(added) else -1
        sr = int(removed[0]) if len(removed) else -1
        adj = data.network['adj_mat']
        if kind == 'grow':
            o = orc.grow_log_q(st['zones'], zz, sn, adj, 0.85)
        elif kind == 'shrink':
            o = orc.shrink_log_q(st['zones'], zz, sr, adj, 0.85)
        else:
            o = orc.swap_log_q(st['zones'], zz, sn, sr, adj, 0.85)
        errs['q_' + kind] = max(errs.get('q_' + kind, 0.), rel_err(o[0], lq), rel_err(o[1], lqb))
        moves.append((['grow', 'shrink', 'swap'].index(kind), zz, sn, sr, lq, lqb))
    out['moves'] = np.array(moves, dtype=float)

    bad = {k: v for k, v in errs.items() if v > 1e-11}
    print(f'{name:18s} ll_nosrc={out["ll_nosrc"]:.6f} ll_src={out["ll_src"]:.6f} prior={out["prior"]:.6f} '
          f'max-err={max(errs.values()):.2e}')
    assert not bad, bad

    adj = data.network['adj_mat'].tocsr()
    save = dict(
        features_idx=orc.state_index(data.features), n_states=np.int64(s), states=data.states,
        families=fam_b, adj_indptr=adj.indptr.astype(np.int32), adj_indices=adj.indices.astype(np.int32),
        cost=(data.geo_prior['cost_matrix'] if geo == 'cost_based' else np.zeros((0, 0))), inheritance=np.bool_(inh),
        zones=st['zones'], weights=st['weights'], p_global=st['p_global'], p_zones=st['p_zones'],
        source_idx=np.argmax(src, axis=-1).astype(np.uint8),
        geo_type=np.str_(geo), size_type=np.str_(size), geo_scale=np.float64(geo_scale), counts_prior=np.bool_(cp),
        scale_universal=np.float64(np.nan if scale_u is None else scale_u),
        scale_inheritance=np.float64(np.nan if scale_f is None else scale_f),
        **{'ref_' + k: np.asarray(v) for k, v in out.items()})
    if inh:
        save['p_families'] = st['p_families']
    if cp:
        save['gibbs_counts_global'] = gibbs_g
        save['alpha_global'] = ag
        if inh:
            save['gibbs_counts_fam'] = gibbs_f
            save['alpha_fam'] = af
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', name + '.npz'), **save)


if __name__ == '__main__':
    os.chdir('/tmp')
    only = sys.argv[1:]
    for i, (name, spec) in enumerate(list(CASES.items()) + list(BUNDLED_CASES.items())):
        if not only or name in only:
            run_case(name, spec, seed=11 + i)
    print('golden vectors written to tests/golden/')
