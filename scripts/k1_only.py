"""Run only the K1 likelihood leg of bench.py (short; the command profiled by ncu for the K1 captures)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

torch.cuda.set_device(0)
torch.cuda.set_stream(torch.cuda.Stream())
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
sizes = tuple(int(v) for v in sys.argv[2].split(',')) if len(sys.argv) > 2 else (500, 2000, 20000)
for r in bench.likelihood_leg(torch, 0, n_launch=n, sizes=sizes):
    print(r)
