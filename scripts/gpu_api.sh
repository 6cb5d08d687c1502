python -m pytest tests/test_step_features.py tests/test_sampler_api.py tests/test_sampler_parity.py -m gpu -x -q 2>&1 | tail -3
python -m pytest tests/test_multi_gpu.py -m gpu -q 2>&1 | tail -2
n=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 5 --warmup 3 --seeds 0 --no-likelihood-leg --no-variants --no-weak > gpurun_out/b2.json 2> gpurun_out/b2.err; tail -2 gpurun_out/b2.err
python -c "
import json; d=json.load(open('gpurun_out/b2.json')); print('N', d['n_gpus'], 'value', d['value'], 'e2e', d['e2e']['value'], 'api', d['api_e2e']['shape_d']['value'])"
python bench.py --steps 5 --warmup 3 --seeds 0 --no-likelihood-leg --no-variants --no-cpu-baseline > gpurun_out/b1.json 2> gpurun_out/b1.err; tail -2 gpurun_out/b1.err
python -c "
import json; d=json.load(open('gpurun_out/b1.json')); print('N', d['n_gpus'], 'value', d['value'], 'e2e', d['e2e']['value'], 'api', {k: v['value'] for k, v in d['api_e2e'].items()})"
