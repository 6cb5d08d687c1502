"""CPU: the numpy port of the reference sampler (oracle/sbayes_sampler_oracle.py -- the timed CPU baseline of bench.py
and the distributional checker) against the posterior summaries of the UNMODIFIED reference (tests/golden/run_*.npz,
oracle/make_golden_runs.py).  One chain of two thirds of the reference's run length; bar |z| < 5."""
import numpy as np
import pytest

import sbayes_sampler_oracle as so
from helpers import PRECISION, load_run, run_cases, run_z_scores, summarise_run


# the full-form MC3 / sim_exp2 fixtures (run_G .. run_I) are multi-chain API runs: the GPU suite covers them through
# ZoneMCMC (tests/test_sampler_parity.py); the single-chain port is pinned on the six single-chain scenarios
SINGLE_CHAIN = [n for n in run_cases() if n[4] <= 'F']


@pytest.mark.parametrize('name', SINGLE_CHAIN)
def test_port_matches_reference_summaries(name):
    c = load_run(name)
    g = c['g']
    st = {'zones': g['zones'], 'weights': g['weights'], 'p_global': g['p_global'], 'p_zones': g['p_zones'],
          'p_families': c['p_families'],
          'source': np.eye(c['C'], dtype=bool)[g['source_idx']] if c['sample_source'] else None}
    smp = so.OracleSampler(c['features'], g['states'], c['families'], c['adj'], c['inheritance'], c['sample_source'],
                           c['prior_cfg'], c['operators'], PRECISION, 3, int(g.get('max_m', 30)), 0.85, st, seed=11,
                           cost_mat=c['cost'],
                           gibbs_counts_global=g.get('gibbs_counts_global'), alpha_global=g.get('alpha_global'),
                           gibbs_counts_fam=g.get('gibbs_counts_fam'), alpha_fam=g.get('alpha_fam'))
    n_steps, every = int(g['n_steps']) * 2 // 3, int(g['n_steps']) // int(g['n_samples'])
    if c['N'] > 500:        # sim_exp1's 951 sites: the port does ~200 steps/s; keep the CPU suite within minutes
        n_steps = min(n_steps, 8000)
    zs, ws, lls, prs = [], [], [], []
    for i in range(n_steps):
        smp.step()
        if i % every == 0:
            zs.append(smp.st.zones.copy()); ws.append(smp.st.weights.copy()); lls.append(smp.ll); prs.append(smp.prior)
    names = [str(k) for k in g['op_names']]
    acc = [smp.accepted.get(k, 0) / max(1, smp.accepted.get(k, 0) + smp.rejected.get(k, 0)) for k in names]
    z = run_z_scores(c, [summarise_run(zs, ws, lls, prs, acc, g['planted'], burn=len(zs) // 5)])
    worst = {k: float(np.max(v)) for k, v in z.items()}
    assert max(worst.values()) < 5.0, (name, worst)
