"""Data formats in front of the hot path: the reference's CSV inputs -> the arrays the device
needs.  Mirrors `sbayes/load_data.py:Data`, `util.read_features_from_csv` / `encode_states` (util.py:289-414),
`preprocessing.compute_network` (:101-174) and the count / cost readers (:519-703), with two differences:

* the encoder emits the `uint8` state index (N, F) the kernels consume (255 = NA) directly; the one-hot bool
  (N, F, S) array of the reference API is derived from it once, instead of the other way round;
* `Data.load_geo_cost_matrix()` is part of the normal loading sequence (the reference CLI forgets it, SURVEY quirk list).

Geodesic distances (`data.CRS`) need cartopy / pyproj, which this image does not have: a CRS raises ImportError as in
the reference (`preprocessing.py:120-126`)."""
import logging

import numpy as np
import pandas as pd
from scipy import spatial
from scipy.sparse import csr_matrix

NA = 255


def _strip(s):
    return s if pd.isna(s) else str.strip(s)


def _strip_frame(df):
    return df.map(_strip) if hasattr(df, 'map') else df.applymap(_strip)


def one_hot(idx, n_states):
    """uint8 state index (N, F), 255 = NA -> bool (N, F, S) with an all-zero row for NA (model.py:112)."""
    out = np.zeros(idx.shape + (n_states,), dtype=bool)
    n, f = np.nonzero(idx != NA)
    out[n, f, idx[n, f]] = True
    return out


def encode_states_uint8(features_raw, feature_states):
    """`util.encode_states` (util.py:289-337) with a uint8 result: state index per (site, feature), 255 = NA.
    Returns (idx, state_names, applicable_states, na_number)."""
    n_states, n_features = feature_states.shape
    n_sites = features_raw.shape[0]
    if n_states > 32:
        raise ValueError('more than 32 states per feature are not supported by the device layout')
    idx = np.full((n_sites, n_features), NA, dtype=np.uint8)
    applicable = np.zeros((n_features, n_states), dtype=bool)
    state_names = {'external': [], 'internal': []}
    na_number = 0
    for f, f_name in enumerate(feature_states.columns):
        f_states = feature_states[f_name]
        applicable[f] = ~f_states.isna()
        s_ext = f_states.dropna().to_list()
        state_names['external'].append(s_ext)
        state_names['internal'].append(list(range(len(s_ext))))
        code = {s: i for i, s in enumerate(s_ext)}
        col = features_raw[f_name]
        unknown = set(col.dropna()) - set(s_ext)
        assert not unknown, f'feature {f_name}: states {sorted(unknown)} are not listed in the feature_states file'
        enc = col.map(code)
        ok = ~enc.isna().to_numpy()
        idx[ok, f] = enc[ok].astype(int).to_numpy()
        na_number += int(np.count_nonzero(~ok))
    return idx, state_names, applicable, na_number


def read_features_from_csv(file, feature_states_file, return_index=False):
    """`util.read_features_from_csv` (util.py:345-414).  Same 9-tuple as the reference; with `return_index` the uint8
    state index is appended."""
    data = _strip_frame(pd.read_csv(file, dtype=str))
    try:
        x, y = data.pop('x'), data.pop('y')
        l_id, name, family = data.pop('id'), data.pop('name'), data.pop('family')
    except KeyError:
        raise KeyError('The csv must contain columns "x", "y", "id","name", "family"')
    feature_states = _strip_frame(pd.read_csv(feature_states_file, dtype=str))
    assert set(feature_states.columns) == set(data.columns)
    n_sites, n_features = data.shape
    locations = np.column_stack([x.astype(float).to_numpy(), y.astype(float).to_numpy()])
    sites = {'locations': locations, 'id': list(range(n_sites)), 'cz': None, 'names': name}
    site_names = {'external': l_id, 'internal': list(range(n_sites))}
    idx, state_names, applicable, na_number = encode_states_uint8(data, feature_states)
    features = one_hot(idx, feature_states.shape[0])
    feature_names = {'external': feature_states.columns.to_numpy(), 'internal': list(range(n_features))}
    fam_ordered = np.unique(family.dropna()).tolist()
    families = np.zeros((len(fam_ordered), n_sites), dtype=int)
    for k, fam in enumerate(fam_ordered):
        families[k, np.where(family == fam)] = 1
    family_names = {'external': fam_ordered, 'internal': list(range(len(fam_ordered)))}
    log = f"{n_sites} sites with {n_features} features read from {file}. {na_number} NA value(s) found."
    out = (sites, site_names, features, feature_names, state_names, applicable, families, family_names, log)
    return out + (idx,) if return_index else out


def compute_delaunay(locations):
    """util.py:158-175: Delaunay neighbours as a CSR adjacency matrix (qhull options as in the reference)."""
    n = len(locations)
    tri = spatial.Delaunay(locations, qhull_options="QJ Pp")
    indptr, indices = tri.vertex_neighbor_vertices
    return csr_matrix((np.ones_like(indices), indices, indptr), shape=(n, n))


class compute_network:
    """`preprocessing.compute_network` (:101-174): a dict-like network with the keys the sampler reads."""

    # above this number of sites the dense N x N distance matrix (8 N^2 bytes: 3.2 GB at N = 20000) is not built unless
    # asked for; the geo prior then works on the site coordinates (implicit Euclidean cost, `sb_upload_locations`)
    DENSE_LIMIT = 8192

    def __init__(self, sites, subset=None, crs=None, dense_distances=None):
        if crs is not None:
            raise ImportError("Using a coordinate reference system (crs) requires the cartopy / pyproj libraries, "
                              "which are not available in this build")
        if subset is None:
            vertices, locations, self.names = sites['id'], np.asarray(sites['locations']), sites['names']
        else:
            sub = np.nonzero(subset)[0]
            vertices, locations = list(range(len(sub))), np.asarray(sites['locations'])[sub]
            self.names = [sites['names'][i] for i in sub]
        adj = compute_delaunay(locations)
        coo = adj.tocoo()
        self.vertices = vertices
        self.edges = np.column_stack((coo.row, coo.col))
        self.locations = locations
        self.adj_mat = adj.tocsr()
        self.n = len(vertices)
        self.m = self.edges.shape[0]
        loc = np.asarray(sites['locations'])
        if dense_distances is None:
            dense_distances = len(loc) <= self.DENSE_LIMIT
        # (row by row: the broadcast form needs a second N x N x 2 temporary)
        self.dist_mat = (np.stack([np.sqrt(((loc - p) ** 2).sum(axis=1)) for p in loc]) if dense_distances and len(loc)
                         else None)

    def __getitem__(self, key):
        return getattr(self, key)

    def get(self, key, default=None):
        return getattr(self, key, default)


def read_feature_occurrence_from_csv(file, feature_states_file):
    """util.py:495-545: aggregated state counts per feature, aligned with the feature_states file."""
    raw = pd.read_csv(file, index_col='feature')
    feature_states = pd.read_csv(feature_states_file, dtype=str)
    n_states, n_features = feature_states.shape
    assert set(raw.index) == set(feature_states.columns)
    raw = raw.fillna(0.)
    feature_names = {'external': feature_states.columns.to_list(), 'internal': list(range(n_features))}
    state_names = {'external': [[] for _ in range(n_features)], 'internal': [[] for _ in range(n_features)]}
    counts = np.zeros((n_features, n_states))
    for f, f_name in enumerate(feature_states.columns):
        for s in range(n_states):
            s_name = feature_states[f_name][s]
            if pd.isnull(s_name):
                continue
            counts[f, s] = raw.loc[f_name, s_name]
            state_names['external'][f].append(s_name)
            state_names['internal'][f].append(s)
    return counts.astype(int), feature_names, state_names


def _counts_from(file, file_type, feature_states_file):
    if file_type == 'counts_file':
        return read_feature_occurrence_from_csv(file, feature_states_file)
    _, _, features, feature_names, state_names, *_ = read_features_from_csv(file, feature_states_file)
    return np.sum(features, axis=0), feature_names, state_names


def _check_names(feature_names, state_names, feature_names_file, state_names_file, file):
    if len(feature_names['external']) != len(feature_names_file['external']):
        raise ValueError("Different number of features in " + str(file) + " as in features.")
    for f in range(len(feature_names['external'])):
        if feature_names['external'][f] != feature_names_file['external'][f]:
            raise ValueError(f"The external feature {f + 1} in {file} differs from the one used in features.")
    if len(state_names['external']) != len(state_names_file['external']):
        raise ValueError("Different number of features in " + str(file) + " as in features.")
    for f in range(len(state_names['external'])):
        if list(state_names['external'][f]) != list(state_names_file['external'][f]):
            raise ValueError(f"The external category names for feature {f + 1} in {file} differ from those used "
                             "in features.")


def read_universal_counts(feature_names, state_names, file, file_type, feature_states_file):
    """preprocessing.py:519-574."""
    counts, fn, sn = _counts_from(file, file_type, feature_states_file)
    n_states = max(len(s) for s in sn['external'])
    assert counts.shape == (len(fn['external']), n_states)
    if not np.all(np.asarray(counts, dtype=float) % 1 == 0):
        raise ValueError(f"The data in {file} must be count data.")
    _check_names(feature_names, state_names, fn, sn, file)
    return counts, f"Read universal counts from {file}"


def read_inheritance_counts(family_names, feature_names, state_names, files, file_type, feature_states_file):
    """preprocessing.py:595-674: (Fam, F, S) counts; a family without a file keeps zeros (= uniform prior)."""
    n_states = max(len(s) for s in state_names['external'])
    counts_all = np.zeros([len(family_names['external']), len(feature_names['external']), n_states])
    log = str()
    for k, fam in enumerate(family_names['external']):
        if fam not in files:
            log += f"No prior information for {fam}. Uniform prior used instead.\n"
            continue
        counts, fn, sn = _counts_from(files[fam], file_type, feature_states_file)
        if not np.all(np.asarray(counts, dtype=float) % 1 == 0):
            raise ValueError(f"The data in {files[fam]} must be count data.")
        _check_names(feature_names, state_names, fn, sn, files[fam])
        counts_all[k] = counts
        log += f"Read counts for {fam} from {files[fam]}\n"
    return counts_all.astype(int), log


def read_geo_cost_matrix(site_names, file):
    """preprocessing.py:677-703: cost matrix ordered like the sites, symmetrised if necessary."""
    costs = pd.read_csv(file, dtype=str, index_col=0)
    log = f"Geographical cost matrix read from {file}."
    assert set(costs.columns) == set(site_names['external'])
    ordered = costs.loc[list(site_names['external']), list(site_names['external'])]
    m = np.asarray(ordered).astype(float)
    if not np.allclose(m, m.T):
        m = (m + m.T) / 2
        log += (".The cost matrix is not symmetric. It was made symmetric by averaging the original costs along the "
                "upper and lower triangle.")
    return m, log


class Data:
    """`sbayes.load_data.Data` (load_data.py:24-143)."""

    def __init__(self, experiment):
        self.path_results = experiment.path_results
        self.experiment_name = experiment.experiment_name
        self.config = experiment.config
        if self.config['data'].get('CRS') is not None:
            raise ImportError("data.CRS needs pyproj / cartopy, which are not available in this build")
        self.crs = None
        self.sites = self.site_names = self.features = self.features_idx = self.feature_names = None
        self.states = self.state_names = self.families = self.family_names = self.network = None
        self.log_load_features = self.log_load_universal_counts = None
        self.log_load_inheritance_counts = self.log_load_geo_cost_matrix = None
        self.prior_universal, self.prior_inheritance, self.geo_prior = {}, {}, {}
        self.is_simulated = False

    def load_features(self):
        (self.sites, self.site_names, self.features, self.feature_names, self.state_names, self.states, self.families,
         self.family_names, self.log_load_features, self.features_idx) = read_features_from_csv(
            file=self.config['data']['FEATURES'], feature_states_file=self.config['data']['FEATURE_STATES'],
            return_index=True)
        self.network = compute_network(self.sites, crs=self.crs, dense_distances=self.config['data'].get('DENSE_DISTANCES'))

    def load_universal_counts(self):
        cfg = self.config['model']['PRIOR']['universal']
        if cfg['type'] != 'counts':
            return
        counts, self.log_load_universal_counts = read_universal_counts(
            feature_names=self.feature_names, state_names=self.state_names, file=cfg['file'],
            file_type=cfg['file_type'], feature_states_file=self.config['data']['FEATURE_STATES'])
        self.prior_universal = {'counts': counts, 'states': self.states}

    def load_inheritance_counts(self):
        if not self.config['model']['INHERITANCE']:
            return
        cfg = self.config['model']['PRIOR']['inheritance']
        if cfg['type'] != 'counts':
            return
        counts, self.log_load_inheritance_counts = read_inheritance_counts(
            family_names=self.family_names, feature_names=self.feature_names, state_names=self.state_names,
            files=cfg['files'], file_type=cfg['file_type'], feature_states_file=self.config['data']['FEATURE_STATES'])
        self.prior_inheritance = {'counts': counts, 'states': self.state_names['internal']}

    def load_geo_cost_matrix(self):
        geo = self.config['model']['PRIOR']['geo']
        if geo['type'] != 'cost_based':
            return
        if 'file' not in geo:
            cost = self.network['dist_mat']
            if cost is None:
                # no dense matrix (data.DENSE_DISTANCES = false, or more than compute_network.DENSE_LIMIT sites): the
                # device derives the same Euclidean costs from the coordinates
                self.geo_prior = {'cost_matrix': None, 'locations': np.asarray(self.network['locations'], float)}
                return
        else:
            cost, self.log_load_geo_cost_matrix = read_geo_cost_matrix(site_names=self.site_names, file=geo['file'])
        self.geo_prior = {'cost_matrix': cost}

    def log_loading(self):
        logging.basicConfig(format='%(message)s', filename=self.path_results / 'experiment.log', level=logging.DEBUG)
        logging.info("\n")
        logging.info("DATA IMPORT")
        logging.info("##########################################")
        for msg in (self.log_load_features, self.log_load_universal_counts, self.log_load_inheritance_counts,
                    self.log_load_geo_cost_matrix):
            logging.info(msg)
