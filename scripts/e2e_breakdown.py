"""Where the e2e step of bench.py spends its time: upload of all chain states (sb_set_states, incl. the recount),
the persistent launch, the read-back of one chain's state and of all (ll, prior)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

nc = 64
data = bench.d_data()
eng = Engine(bench.state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=7, n_chains=nc,
             inheritance=True, sample_source=True)
states = [bench.d_state(data, 100 + c) for c in range(nc)]
host = {k: torch.from_numpy(np.ascontiguousarray(np.stack(v))).pin_memory() for k, v in dict(
    zones=[s['zones'].astype(np.uint8) for s in states], weights=[s['weights'] for s in states],
    p_global=[s['p_global'][0] for s in states], p_zones=[s['p_zones'] for s in states],
    p_families=[s['p_families'] for s in states],
    source=[np.argmax(s['source'], -1).astype(np.uint8) for s in states]).items()}


def upload():
    eng.set_states(0, host['zones'].numpy(), host['weights'].numpy(), host['p_global'].numpy(),
                   host['p_zones'].numpy(), host['p_families'].numpy(), host['source'].numpy())


upload(); eng.eval()
eng.set_operators(bench.default_ops(), bench.PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=1)
eng.run(2000); eng.sync()
T = {'upload (call returns)': [], 'upload (stream drained)': [], 'run 2000 steps': [], 'get_state(0)': [],
     'chain_scalars': []}
for _ in range(5):
    t0 = time.perf_counter(); upload(); t1 = time.perf_counter(); eng.sync(); t2 = time.perf_counter()
    eng.run(2000); eng.sync(); t3 = time.perf_counter()
    eng.get_state(0, one_hot_source=False); t4 = time.perf_counter()
    eng.chain_scalars(); t5 = time.perf_counter()
    for k, v in zip(T, (t1 - t0, t2 - t0, t3 - t2, t4 - t3, t5 - t4)):
        T[k].append(v)
for k, v in T.items():
    print(f'{k:26s} {1e3 * np.median(v):8.3f} ms')
