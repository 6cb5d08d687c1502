"""api_e2e leg alone, a few repetitions, with the duration of every host phase of generate_samples (SBAYES_B200_TRACE)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

torch.cuda.set_device(0)
data = bench.d_data()
ring = int(sys.argv[1]) if len(sys.argv) > 1 else None
from contact_zones_b200.sampling import ZoneMCMC  # noqa: E402
if ring:
    ZoneMCMC.RING_BYTES_PER_LAUNCH = ring << 20
for rep in range(4):
    r = bench.api_leg(torch, data, 64, 100000, 100, 0)
    print('ring MB', ring, 'rep', rep, 'value', round(r['value'] / 1e6, 2), 'sampling_time', round(r['sampling_time_s'], 4))
