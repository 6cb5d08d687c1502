"""Worker of tests/test_multi_gpu.py: one rank of `ZoneMCMC(n_chains=8, mc3=True)` under torch.distributed (NCCL, one
process per GPU).  Rank 0 writes the logged statistics to the .npz named on the command line."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'oracle'), os.path.join(ROOT, 'tests')):
    sys.path.insert(0, p)


def run(out_path, distributed):
    from helpers import PRECISION, case_data, case_model_config, load_case
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import ZoneMCMC
    if distributed:
        import torch
        import torch.distributed as dist
        local = int(os.environ['LOCAL_RANK'])
        torch.cuda.set_device(local)
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    c = load_case('C_samerica_like')
    data = case_data(c)
    model = Model(data, case_model_config(c))
    ops = {'shrink_zone': 0.04, 'grow_zone': 0.04, 'swap_zone': 0.02, 'gibbs_sample_weights': 0.4,
           'gibbs_sample_p_global': 0.1, 'gibbs_sample_p_zones': 0.3, 'gibbs_sample_p_families': 0.1}
    smp = ZoneMCMC(data=data, model=model, n_chains=8, operators=ops, var_proposal=PRECISION, p_grow_connected=0.85,
                   initial_size=5, mc3=True, swap_period=60, chain_swaps=3, seed=77)
    smp.generate_samples(3000, 30)
    st = smp.statistics
    rank = int(os.environ.get('RANK', 0))
    if rank == 0:
        names = sorted(st['accept_operator'])
        np.savez(out_path, zones=np.stack(st['sample_zones']), weights=np.stack(st['sample_weights']),
                 p_global=np.stack(st['sample_p_global']), p_zones=np.stack(st['sample_p_zones']),
                 p_families=np.stack(st['sample_p_families']), ll=np.array(st['sample_likelihood']),
                 prior=np.array(st['sample_prior']), chain_idx=np.array(smp.chain_idx),
                 swaps=np.array([st['n_swaps'], st['accepted_swaps']]),
                 acc=np.array([st['accept_operator'][k] for k in names]), rej=np.array([st['reject_operator'][k] for k in names]),
                 last_zones=st['last_sample'].zones, last_source=st['last_sample'].source,
                 n_local=np.int64(smp.shard.n_local), world=np.int64(smp.shard.world))
    smp.close()
    if distributed:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    run(sys.argv[1], distributed='RANK' in os.environ)
