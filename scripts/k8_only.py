"""The device-resident sampler leg of bench.py alone (one seed, short): the command profiled by ncu for the step kernel.
usage: python scripts/k8_only.py <chains_total> [timed_steps]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

chains = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
torch.cuda.set_device(0)
torch.cuda.set_stream(torch.cuda.Stream())
stream = torch.cuda.current_stream().cuda_stream
args = type('A', (), dict(mcmc_steps=5000, chain_warm_steps=2000))()
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
leg = bench.sampler_leg(torch, args, bench.d_data(), chains, 0, stream, torch.cuda.synchronize, flush, steps, 1)
print({k: leg[k] for k in ('kernel_ms', 'step_ms', 'launches', 'drift')}, 'us/step/chain', 1e3 * leg['kernel_ms'] / 5000)
