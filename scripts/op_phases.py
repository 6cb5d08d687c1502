"""Phase timing of the step kernel (needs the -DSB_PROF build: csrc/libsbayes_b200_prof.so via SBAYES_B200_LIB)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import bench  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 500
data = bench.d_data()
eng = Engine(bench.state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=7, n_chains=64,
             inheritance=True, sample_source=True)
for c in range(64):
    s = bench.d_state(data, 100 + c)
    eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], s['p_families'], np.argmax(s['source'], -1))
eng.eval()
buf = (C.c_ulonglong * 16)()
names = ['gibbs:phase0+1', 'gibbs:sync', 'gibbs:phase2', 'gibbs:reduce', 'step:op-select', 'step:operator', 'step:tail-sync', 'w:all', 'z:nbr', 'z:count', 'z:proposal', 'z:delta', 'z:prelude', 'z:decide']
for k in ('swap_zone', 'grow_zone', 'shrink_zone'):
    eng.set_operators({k: 1.0}, bench.PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=1)
    eng.run(20); eng.sync()
    eng.lib.sb_prof_read(buf, 1)
    eng.run(n_steps); eng.sync()
    eng.lib.sb_prof_read(buf, 1)
    print(k, ' '.join(f'{n}={buf[i] / n_steps:.0f}' for i, n in enumerate(names)), flush=True)
