"""Fixed cost of a step of K8: gibbs_sample_p_global with (almost) no feature selected -- the step then consists of
the loop overhead, the operator dispatch and the per-row subset draw only -- against the normal operator, and
two-operator mixes that show how much consecutive barrier-free Gibbs steps overlap."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import bench  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
nc = 64
data = bench.d_data()
eng = Engine(bench.state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=7, n_chains=nc,
             inheritance=True, sample_source=True)
for c in range(nc):
    s = bench.d_state(data, 100 + c)
    eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], s['p_families'], np.argmax(s['source'], -1))
eng.eval()
cases = [('p_global frac 1e-9', {'gibbs_sample_p_global': 1.0}, 1e-9),
         ('p_global frac 0.4', {'gibbs_sample_p_global': 1.0}, 0.4),
         ('p_global frac 1.0', {'gibbs_sample_p_global': 1.0}, 1.0),
         ('p_zones', {'gibbs_sample_p_zones': 1.0}, 0.4),
         ('weights', {'gibbs_sample_weights': 1.0}, 0.4),
         ('p_zones .5 + weights .5', {'gibbs_sample_p_zones': .5, 'gibbs_sample_weights': .5}, 0.4),
         ('p_zones .5 + p_global(1e-9) .5', {'gibbs_sample_p_zones': .5, 'gibbs_sample_p_global': .5}, 1e-9),
         ('weights .5 + p_global(1e-9) .5', {'gibbs_sample_weights': .5, 'gibbs_sample_p_global': .5}, 1e-9)]
for name, ops, frac in cases:
    eng.set_operators(ops, bench.PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=1,
                      gibbs_feature_fraction=frac)
    eng.run(20); eng.sync()
    t0 = time.perf_counter(); eng.run(n_steps); eng.sync(); dt = time.perf_counter() - t0
    print(f'{name:34s} {1e6 * dt / n_steps:9.2f} us/step', flush=True)
