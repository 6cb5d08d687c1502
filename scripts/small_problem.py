"""The drop-in API on a Balkan-sized problem (28 sites x 47 features x 3 states, 3 zones, 5 families, 4 MC3 chains): the
per-step latency floor of small data, by team size (SBAYES_B200_TEAM)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from contact_zones_b200 import synthetic  # noqa: E402

torch.cuda.set_device(0)
small = synthetic.make_data(28, 47, 3, n_families=5, seed=4, with_dist=False, ragged_states=True, planted_zones=3, planted_size=5)
cfg = dict(bench.MODEL_CFG, N_AREAS=3, MAX_M=10)
for rep in range(2):
    r = bench.api_leg(torch, small, 4, 1000000, 1000, 0, cfg)
    print('team', os.environ.get('SBAYES_B200_TEAM', 'auto'), 'us per proposal per chain', round(r['us_per_proposal_per_chain'], 3), 'value', round(r['value'] / 1e6, 3))
