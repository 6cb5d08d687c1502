"""Per-operator cost of the step kernel K8 on shape D with SAMPLE_SOURCE = false (Metropolis alter_* operators)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
nc = 64
data = bench.d_data()
eng = Engine(bench.state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=7, n_chains=nc,
             inheritance=True, sample_source=False)
for c in range(nc):
    s = bench.d_state(data, 100 + c)
    eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], s['p_families'], None)
eng.eval()
mixes = {'default': {'shrink_zone': .02, 'grow_zone': .02, 'swap_zone': .01, 'alter_weights': .4, 'alter_p_global': .05,
                     'alter_p_zones': .4, 'alter_p_families': .1}}
for k in ('alter_weights', 'alter_p_global', 'alter_p_zones', 'alter_p_families', 'grow_zone', 'shrink_zone', 'swap_zone'):
    mixes[k] = {k: 1.0}
for name, ops in mixes.items():
    eng.set_operators(ops, bench.PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=1)
    eng.run(20); eng.sync()
    t0 = time.perf_counter(); eng.run(n_steps); eng.sync(); dt = time.perf_counter() - t0
    acc, rej = eng.op_stats()
    print(f'{name:26s} {1e6 * dt / n_steps:9.2f} us/step   (acc {acc.sum()} rej {rej.sum()})', flush=True)
