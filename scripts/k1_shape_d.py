"""K1 / K1s / K4 on the headline shape D (bench.shape_d_kernel_leg) -- the command profiled by ncu for those kernels."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

torch.cuda.set_device(0)
torch.cuda.set_stream(torch.cuda.Stream())
peak, _ = bench.measured_peak()
out = bench.shape_d_kernel_leg(torch, bench.d_data(), 0, torch.cuda.current_stream().cuda_stream, peak)
for k, v in out.items():
    print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a in ('us_per_launch', 'us_min', 'frac', 'achieved')})
# one more launch of each kernel, back to back, after everything else: the window an ncu capture of this command takes
# (the matched launches before it: 64 one-chain recounts of sb_set_state + 3 x 45 warm-up / timed launches)
from contact_zones_b200 import synthetic  # noqa: E402
from contact_zones_b200.engine import Engine  # noqa: E402

d, data = bench.D_SHAPE, bench.d_data()
eng = Engine(bench.state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=d['Z'], n_chains=64,
             inheritance=True, sample_source=True, device=0)
sts = [synthetic.random_state(data, d['Z'], 20, True, True, seed=300 + i) for i in range(8)]
for c in range(64):
    s = sts[c % 8]
    eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], s['p_families'], s['source'])
eng.sync()
eng.eval_likelihood_async(use_source=False)
eng.eval_likelihood_async(use_source=True)
eng.recount_async()
eng.sync()
eng.close()
