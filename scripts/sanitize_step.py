"""Small sampler runs for compute-sanitizer (memcheck / racecheck): the step kernel with teams of 1 and 2 CTAs, both
SAMPLE_SOURCE modes, zone moves + every parameter operator, events + snapshot ring, then the evaluation kernels.
usage: SBAYES_B200_TEAM=<1|2> python scripts/sanitize_step.py <n_steps>"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from contact_zones_b200 import synthetic  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
PREC = {'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20}
data = synthetic.make_data(150, 40, 4, n_families=3, seed=5, with_dist=True, planted_zones=2, planted_size=8)
for src in (True, False):
    eng = Engine(data.features, data.states, data.families, data.network['adj_mat'], data.geo_prior['cost_matrix'],
                 n_zones=2, n_chains=3, inheritance=True, sample_source=src)
    eng.set_priors(size_type='uniform', geo_type='cost_based', geo_scale=100.0)
    for c in range(3):
        st = synthetic.random_state(data, 2, 6, True, src, seed=10 + c)
        eng.set_state(c, st['zones'], st['weights'], st['p_global'], st['p_zones'], st['p_families'], st['source'])
    ll0, pr0, _ = eng.eval()
    kind = 'gibbs_sample' if src else 'alter'
    ops = {'shrink_zone': .1, 'grow_zone': .1, 'swap_zone': .1, f'{kind}_weights': .2, f'{kind}_p_global': .15,
           f'{kind}_p_zones': .2, f'{kind}_p_families': .15}
    if src:
        ops['gibbs_sample_sources'] = 0.02
    eng.set_operators(ops, PREC, min_size=3, max_size=12, p_grow_connected=0.85, seed=3)
    eng.run(n_steps)
    eng.run_events(n_steps, [n_steps // 3, n_steps // 2, n_steps], [0, -1, 1], ring_chain=-1)
    ll_ev, pr_ev = eng.event_scalars(3)
    snap = eng.read_snapshot(1, 2)
    ll, pr = eng.chain_scalars()
    ll2, pr2, _ = eng.eval()
    assert np.allclose(ll, ll2, rtol=1e-9) and np.allclose(pr, pr2, rtol=1e-9, atol=1e-9), (ll, ll2, pr, pr2)
    assert np.array_equal(ll_ev[2], ll) and snap['zones'].sum() >= 6
    q = eng.move_log_q(0, 'grow', 0, int(np.flatnonzero(~eng.get_state(0)['zones'].any(0))[0]), -1, 0.85)
    if src:
        eng.counts(0, recount=True)
    print(f'sample_source={src}: team {os.environ.get("SBAYES_B200_TEAM", "auto")}, {2 * n_steps} steps x 3 chains ok, '
          f'll {ll[0]:.3f}, log_q {q[0]:.4f}')
    eng.close()
