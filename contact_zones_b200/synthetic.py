"""Synthetic `data` objects and random sampler states of the shapes named in BASELINE.json / SURVEY.md §8(d).

The `data` duck type is the one the reference's hot path consumes (`zone_sampling.py:128-135,151`,
`model.py:106-107,473,542,583,659,790,1009`): `.features` bool (N,F,S), `.families` bool (Fam,N), `.states`
bool (F,S), `.network{'adj_mat','locations','dist_mat'}`, `.prior_universal`, `.prior_inheritance`,
`.geo_prior`, `.state_names`, `.is_simulated`.  Host-side numpy only (input producer, not on the hot path).
"""
import types

import numpy as np
from scipy import spatial
from scipy.sparse import csr_matrix


def delaunay_csr(locations):
    """Delaunay adjacency as CSR -- same construction as the reference's input producer
    (`util.py:158-175`: `spatial.Delaunay(qhull_options="QJ Pp").vertex_neighbor_vertices`)."""
    n = len(locations)
    tri = spatial.Delaunay(locations, qhull_options="QJ Pp")
    indptr, indices = tri.vertex_neighbor_vertices
    return csr_matrix((np.ones_like(indices), indices, indptr), shape=(n, n))


def make_data(n_sites, n_features, n_states, n_families=0, na_frac=0.02, seed=0, ragged_states=False,
              with_dist=True, counts_prior=False, planted_zones=0, planted_size=20, planted_conc=0.5):
    """Synthetic data set of SURVEY.md §8(d): uniform random states, `na_frac` NA cells, random family
    assignment (-1 = none), uniform locations on [0,1000]^2 with Delaunay adjacency.

    planted_zones > 0 draws the members of that many random connected groups from their own
    Dirichlet(planted_conc) state distribution so that area moves are accepted at realistic rates."""
    rng = np.random.default_rng(seed)
    n_states_f = np.full(n_features, n_states)
    if ragged_states and n_states > 2:
        n_states_f = rng.integers(2, n_states + 1, size=n_features)
    states = np.arange(n_states)[None, :] < n_states_f[:, None]
    x = (rng.random((n_sites, n_features)) * n_states_f[None, :]).astype(np.int64)
    locations = rng.random((n_sites, 2)) * 1000.
    adj = delaunay_csr(locations) if n_sites >= 4 else csr_matrix(np.ones((n_sites, n_sites)) - np.eye(n_sites))

    planted = np.zeros((planted_zones, n_sites), dtype=bool)
    taken = np.zeros(n_sites, dtype=bool)
    for z in range(planted_zones):
        seed_site = rng.choice(np.flatnonzero(~taken))
        members = [seed_site]
        taken[seed_site] = True
        while len(members) < planted_size:
            cand = np.flatnonzero(np.asarray(adj[members].sum(axis=0)).ravel().astype(bool) & ~taken)
            if len(cand) == 0:
                break
            s = rng.choice(cand)
            members.append(s)
            taken[s] = True
        planted[z, members] = True
        for f in range(n_features):
            p = rng.dirichlet(np.full(n_states_f[f], planted_conc))
            x[members, f] = rng.choice(n_states_f[f], size=len(members), p=p)

    features = np.zeros((n_sites, n_features, n_states), dtype=bool)
    nn, ff = np.meshgrid(np.arange(n_sites), np.arange(n_features), indexing='ij')
    features[nn, ff, x] = True
    if na_frac > 0:
        features[rng.random((n_sites, n_features)) < na_frac] = False

    if n_families > 0:
        fam_id = rng.integers(-1, n_families, size=n_sites)
        # every family gets at least one member
        fam_id[rng.permutation(n_sites)[:n_families]] = np.arange(n_families)
        families = fam_id[None, :] == np.arange(n_families)[:, None]
    else:
        families = np.zeros((0, n_sites), dtype=bool)

    dist = None
    if with_dist:
        d = locations[:, None, :] - locations[None, :, :]
        dist = np.sqrt((d ** 2).sum(-1))

    data = types.SimpleNamespace(
        features=features, families=families, states=states,
        network={'adj_mat': adj, 'locations': locations, 'dist_mat': dist, 'n': n_sites},
        prior_universal={}, prior_inheritance={}, geo_prior={'cost_matrix': dist},
        state_names={'internal': [list(range(k)) for k in n_states_f],
                     'external': [[f's{j}' for j in range(k)] for k in n_states_f]},
        feature_names={'internal': list(range(n_features)), 'external': [f'f{i}' for i in range(n_features)]},
        family_names={'internal': list(range(n_families)), 'external': [f'fam{i}' for i in range(n_families)]},
        site_names={'internal': list(range(n_sites)), 'external': [f'site{i}' for i in range(n_sites)],
                    'id': list(range(n_sites))},
        is_simulated=False, planted_zones=planted, crs=None)
    if counts_prior:
        data.prior_universal = {'counts': rng.integers(0, 30, size=(n_features, n_states)) * states}
        if n_families > 0:
            data.prior_inheritance = {
                'counts': rng.integers(0, 20, size=(n_families, n_features, n_states)) * states[None]}
    return data


def random_zones(data, n_zones, size, rng):
    """Disjoint connected zones grown by a random neighbour walk (like `zone_sampling.py:992-1031`)."""
    adj = data.network['adj_mat']
    n = data.features.shape[0]
    zones = np.zeros((n_zones, n), dtype=bool)
    occupied = np.zeros(n, dtype=bool)
    for z in range(n_zones):
        for _attempt in range(200):
            zone = np.zeros(n, dtype=bool)
            occ = occupied.copy()
            s = rng.choice(np.flatnonzero(~occ))
            zone[s] = occ[s] = True
            ok = True
            for _ in range(size - 1):
                nb = np.flatnonzero(np.logical_and(adj.dot(zone), ~occ))
                if len(nb) == 0:
                    ok = False
                    break
                s = rng.choice(nb)
                zone[s] = occ[s] = True
            if ok:
                zones[z] = zone
                occupied = occ
                break
        else:
            raise ValueError('could not place zones')
    return zones


def random_state(data, n_zones, zone_size, inheritance, with_source, seed=0, zones=None):
    """A random sampler state (dict of the `Sample` arrays, `zone_sampling.py:68-75`)."""
    rng = np.random.default_rng(seed)
    n, f, s = data.features.shape
    c = 3 if inheritance else 2
    states = data.states

    def dirichlet_rows(lead):
        out = np.zeros(lead + (f, s))
        for i in range(f):
            k = int(states[i].sum())
            out[..., i, states[i]] = rng.dirichlet(np.ones(k), size=lead)
        return out

    st = {
        'zones': random_zones(data, n_zones, zone_size, rng) if zones is None else np.asarray(zones, dtype=bool),
        'weights': rng.dirichlet(np.ones(c), size=f),
        'p_global': dirichlet_rows((1,)),
        'p_zones': dirichlet_rows((n_zones,)),
        'p_families': dirichlet_rows((data.families.shape[0],)) if inheritance else None,
        'source': None,
    }
    if with_source:
        has = np.stack([np.ones(n, bool), st['zones'].any(0)] +
                       ([np.asarray(data.families, bool).any(0)] if inheritance else []), axis=1)
        w = st['weights'][None] * has[:, None, :]
        w = w / w.sum(-1, keepdims=True)
        u = rng.random((n, f, 1))
        src = np.argmax(u < np.cumsum(w, -1), axis=-1)
        st['source'] = np.eye(c, dtype=bool)[src]
    return st
