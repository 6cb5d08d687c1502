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
