"""TEST INFRASTRUCTURE ONLY -- step rate of the numpy port (oracle/sbayes_sampler_oracle.py, the timed CPU arm of
bench.py on the GPU box, where /root/reference does not exist) next to the UNMODIFIED reference's
`ZoneMCMC.generate_samples` on the same data, state, operator mix and core: the port must not be slower than the
reference, or the GPU / CPU ratio of bench.py would be inflated.

Run in the build container:   python oracle/time_vs_reference.py > profiles/r02_port_vs_reference.txt
Shapes: D = synthetic 2000 x 200 x 8, 7 zones, 20 families (BASELINE.json configs[3]); A = 951 x 35 x 4, 1 zone, no
families (experiments/simulation size).  SAMPLE_SOURCE, default operator mix (default_config.json:13-20)."""
import contextlib
import io
import os
import random
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import refshim  # noqa: E402
import sbayes_sampler_oracle as so  # noqa: E402
from contact_zones_b200 import synthetic  # noqa: E402

ref = refshim.modules()
PRECISION = {'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20}
PRIOR = {'geo': {'type': 'uniform'}, 'area_size': {'type': 'none'}, 'weights': {'type': 'uniform'},
         'universal': {'type': 'uniform'}, 'contact': {'type': 'uniform'}, 'inheritance': {'type': 'uniform'}}
SHAPES = {'D (2000 x 200 x 8, 7 zones, 20 families)': dict(N=2000, F=200, S=8, Fam=20, Z=7, steps=300),
          'A (951 x 35 x 4, 1 zone, no families)': dict(N=951, F=35, S=4, Fam=0, Z=1, steps=2000)}


def main():
    os.environ['OMP_NUM_THREADS'] = '1'
    for name, c in SHAPES.items():
        inh = c['Fam'] > 0
        data = synthetic.make_data(c['N'], c['F'], c['S'], n_families=c['Fam'], seed=0, with_dist=False,
                                   planted_zones=c['Z'], planted_size=20)
        st = synthetic.random_state(data, c['Z'], 5, inh, True, seed=100)
        ops = so.default_operators(inh, True)
        prior = {k: v for k, v in PRIOR.items() if inh or k != 'inheritance'}
        model = ref.model.Model(data, {'N_AREAS': c['Z'], 'MIN_M': 3, 'MAX_M': 50, 'INHERITANCE': inh,
                                       'SAMPLE_SOURCE': True, 'PRIOR': prior})
        init = ref.zone_sampling.Sample(zones=st['zones'].copy(), weights=st['weights'].copy(), p_global=st['p_global'].copy(),
                                        p_zones=st['p_zones'].copy(),
                                        p_families=None if st['p_families'] is None else st['p_families'].copy())
        random.seed(1)
        np.random.seed(1)
        smp = ref.zone_sampling.ZoneMCMC(data=data, model=model, n_chains=1, initial_sample=init, operators=ops,
                                         var_proposal=PRECISION, p_grow_connected=0.85, initial_size=5)
        with contextlib.redirect_stdout(io.StringIO()):
            smp.generate_samples(20, 2)                      # untimed: caches, imports
            t0 = time.perf_counter()
            smp.generate_samples(c['steps'], 10)
            t_ref = time.perf_counter() - t0
        # generate_samples re-initialises the chain (one likelihood + prior evaluation): negligible against the steps
        port = so.OracleSampler(data.features, data.states, data.families, data.network['adj_mat'], inh, True, prior, ops,
                                PRECISION, 3, 50, 0.85, st, seed=1)
        port.run(20)
        t0 = time.perf_counter()
        port.run(c['steps'])
        t_port = time.perf_counter() - t0
        print(f'shape {name}: {c["steps"]} steps, 1 chain, 1 core')
        print(f'  unmodified reference  {c["steps"] / t_ref:8.1f} steps/s')
        print(f'  numpy port (CPU arm)  {c["steps"] / t_port:8.1f} steps/s   port / reference = {t_ref / t_port:.2f}')


if __name__ == '__main__':
    main()
