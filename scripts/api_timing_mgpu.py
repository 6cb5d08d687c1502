"""api_e2e leg under torchrun (one rank per GPU), with the host phases of generate_samples traced on rank 0."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402

rank, local = int(os.environ['RANK']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
if rank == 0:
    os.environ['SBAYES_B200_TRACE'] = '1'
data = bench.d_data()
for rep in range(3):
    dist.barrier()
    t0 = time.time()
    r = bench.api_leg(torch, data, 64, 100000, 100, 0)
    if rank == 0:
        print('rep', rep, 'value', round(r['value'] / 1e6, 2), 'sampling_time', round(r['sampling_time_s'], 4), 'whole call', round(time.time() - t0, 3))
dist.destroy_process_group()
