"""Host-side timeline of bench.py's pipelined e2e leg (two contexts alternating): how long every call of a step blocks."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

torch.cuda.set_device(0)
nc, steps = 64, 5000
data = bench.d_data()
smp = bench.make_sampler(data, nc, 0, steps)
eng0 = smp._initialise_chains()
eng0.run(4000)                                        # warmed states, as in bench.py
host = bench.host_states([eng0.get_state(c, one_hot_source=False) for c in range(nc)])
smp.close()
engs = [Engine(bench.state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=7, n_chains=nc,
               inheritance=True, sample_source=True, device=0) for _ in range(2)]


def upload(e):
    e.set_states(0, host['zones'].numpy(), host['weights'].numpy(), host['p_global'].numpy(), host['p_zones'].numpy(),
                 host['p_families'].numpy(), host['source'].numpy())


def read(e):
    e.get_state(0, one_hot_source=False)
    e.chain_scalars()


for e in engs:
    upload(e); e.eval()
    e.set_operators(bench.default_ops(), bench.PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=0)
    e.run(steps); read(e)
torch.cuda.synchronize()
rows = []
t_start = time.perf_counter()
for k in range(8):
    e = engs[k & 1]
    t0 = time.perf_counter(); upload(e)
    t1 = time.perf_counter(); e.run(steps)
    t2 = time.perf_counter()
    if k >= 1:
        read(engs[(k - 1) & 1])
    t3 = time.perf_counter()
    rows.append((t0 - t_start, t1 - t0, t2 - t1, t3 - t2))
torch.cuda.synchronize()
print('step  start_ms  upload_ms  run_call_ms  read_prev_ms')
for k, r in enumerate(rows):
    print(k, ' '.join(f'{1e3 * v:9.3f}' for v in r))
print('per step', 1e3 * (time.perf_counter() - t_start) / 8, 'ms')
# the same with one context, serial
e = engs[0]
t0 = time.perf_counter()
for k in range(4):
    upload(e); e.run(steps); read(e)
print('serial per step', 1e3 * (time.perf_counter() - t0) / 4, 'ms')
# does a concurrent upload into the OTHER context slow the persistent launch down?
for label, do_up in (('launch alone', False), ('launch + concurrent upload', True), ('launch alone', False), ('launch + concurrent upload', True)):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    engs[0].run(steps)
    if do_up:
        upload(engs[1])
    t1 = time.perf_counter()
    engs[0].sync()
    t2 = time.perf_counter()
    engs[1].sync()
    print(f'{label:30s} host calls {1e3 * (t1 - t0):7.3f} ms, launch done after {1e3 * (t2 - t0):7.3f} ms')
# when does the device side of an upload into context 1 finish while context 0's launch is running?
for rep in range(2):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    engs[0].run(steps)
    upload(engs[1])
    t1 = time.perf_counter()
    engs[1].sync()
    t2 = time.perf_counter()
    engs[0].sync()
    t3 = time.perf_counter()
    print(f'upload call {1e3 * (t1 - t0):7.3f} ms, its stream drained after {1e3 * (t2 - t0):7.3f} ms, launch done after {1e3 * (t3 - t0):7.3f} ms')
torch.cuda.synchronize()
t0 = time.perf_counter(); upload(engs[1]); t1 = time.perf_counter(); engs[1].sync(); t2 = time.perf_counter()
print(f'idle GPU: upload call {1e3 * (t1 - t0):7.3f} ms, drained after {1e3 * (t2 - t0):7.3f} ms')
