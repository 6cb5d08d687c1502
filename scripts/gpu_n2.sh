python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
tail -3 gpurun_out/bench_n2.err
python -c "
import json;d=json.load(open('gpurun_out/bench_n2.json'));print({k:d[k] for k in ('value','n_gpus','ms_per_step','us_per_proposal_per_chain','scaling')});print(d['e2e']);print(d['config'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 | cut -c1-300
cat > /tmp/mc3_nccl.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), 'tests')); sys.path.insert(0, os.path.join(os.getcwd(), 'oracle'))
import numpy as np, torch, torch.distributed as dist
rank = int(os.environ['RANK']); torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
from helpers import PRECISION, case_data, case_model_config, load_case
from contact_zones_b200.model import Model
from contact_zones_b200.sampling import ZoneMCMC
c = load_case('C_samerica_like'); data = case_data(c); model = Model(data, case_model_config(c))
ops = {'shrink_zone': 0.02, 'grow_zone': 0.02, 'swap_zone': 0.01, 'gibbs_sample_weights': 0.4, 'gibbs_sample_p_global': 0.05, 'gibbs_sample_p_zones': 0.4, 'gibbs_sample_p_families': 0.1}
smp = ZoneMCMC(data=data, model=model, operators=ops, var_proposal=PRECISION, p_grow_connected=0.85, initial_size=5, n_chains=6, mc3=True, swap_period=50, chain_swaps=2, seed=5)
smp.generate_samples(1000, 10)
st = smp.statistics
print(rank, 'local chains', smp.shard.n_local, 'chain_idx', smp.chain_idx, 'swaps', st['accepted_swaps'], '/', st['n_swaps'], 'll', np.round(st['sample_likelihood'][-3:], 3), 'acc', st['accepted_steps'], flush=True)
dist.destroy_process_group()
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 /tmp/mc3_nccl.py 2>&1 | grep -v Warning | tail -4
