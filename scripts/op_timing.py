"""Per-operator cost of the step kernel K8 on shape D: one launch of n steps with a single operator enabled."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import bench  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
nc = int(sys.argv[2]) if len(sys.argv) > 2 else 64
data = bench.d_data()
eng = Engine(bench.state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=7, n_chains=nc,
             inheritance=True, sample_source=True)
for c in range(nc):
    s = bench.d_state(data, 100 + c)
    eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], s['p_families'], np.argmax(s['source'], -1))
eng.eval()
mixes = {'default': bench.default_ops()}
mixes['gibbs_mix'] = {k: v for k, v in bench.default_ops().items() if k.startswith('gibbs')}
mixes['gibbs_mix+swap'] = dict(mixes['gibbs_mix'], swap_zone=0.05)
mixes['gibbs_mix+shrink'] = dict(mixes['gibbs_mix'], shrink_zone=0.05)
for k in ('gibbs_sample_weights', 'gibbs_sample_p_zones', 'gibbs_sample_p_global', 'gibbs_sample_p_families',
          'grow_zone', 'shrink_zone', 'swap_zone', 'gibbs_sample_sources'):
    mixes[k] = {k: 1.0}
for name, ops in mixes.items():
    eng.set_operators(ops, bench.PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=1)
    n = n_steps if name != 'gibbs_sample_sources' else max(10, n_steps // 20)
    eng.run(20); eng.sync()
    t0 = time.perf_counter(); eng.run(n); eng.sync(); dt = time.perf_counter() - t0
    acc, rej = eng.op_stats()
    print(f'{name:26s} {1e6 * dt / n:9.2f} us/step   (acc {acc.sum()} rej {rej.sum()})', flush=True)
