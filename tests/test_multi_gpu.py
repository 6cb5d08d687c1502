"""GPU (2+ devices): the sharded product itself -- `ZoneMCMC(n_chains=8, mc3=True)` under torchrun with one NCCL rank
per GPU (chains block-partitioned by `ChainShard`, all-gather of (ll, prior), broadcast of the logged chain's snapshot,
rank 0's seed) -- against the same run in one process.  The RNG is keyed by the GLOBAL chain id and the step, the
swap decisions by a shared seeded stream, so the logged samples, `chain_idx` and the operator statistics must agree
bit for bit."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize('world', [2, 4])
def test_sharded_zone_mcmc_equals_single_process(world, built_lib, tmp_path):
    if _n_gpus() < world:
        pytest.skip(f'needs {world} GPUs')
    worker = os.path.join(ROOT, 'tests', 'mgpu_worker.py')
    one, many = str(tmp_path / 'one.npz'), str(tmp_path / 'many.npz')
    env = {k: v for k, v in os.environ.items() if k not in ('RANK', 'LOCAL_RANK', 'WORLD_SIZE')}
    subprocess.run([sys.executable, worker, one], check=True, env=env, timeout=600)
    subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', f'--nproc-per-node={world}',
                    '--master-addr', '127.0.0.1', '--master-port', str(29500 + world), worker, many],
                   check=True, env=env, timeout=900)
    a, b = np.load(one), np.load(many)
    assert int(a['world']) == 1 and int(b['world']) == world and int(b['n_local']) == 8 // world
    for k in a.files:
        if k in ('world', 'n_local'):
            continue
        assert np.array_equal(a[k], b[k]), k
    assert a['swaps'][0] == (3000 // 60) * 3 and a['swaps'][1] > 0
