# strong-scaling runs of bench.py on one box: bash scripts/gpu_scale.sh "2 4 8"   (under gpurun --gpus 8)
for n in ${1:-2 4 8}; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520 + n)) \
    bench.py --gpus $n --steps 10 --warmup 3 --no-likelihood-leg --no-variants > gpurun_out/bench_r02_n$n.json 2> gpurun_out/bench_r02_n$n.err
  echo "== N=$n rc=$?"; tail -c 400 gpurun_out/bench_r02_n$n.err | tail -3
  python -c "
import json;d=json.load(open('gpurun_out/bench_r02_n$n.json'))
print({k:d[k] for k in ('value','n_gpus','ms_per_step','us_per_proposal_per_chain','chains_per_gpu')}); print('weak', d['weak']); print('e2e', d['e2e']['value'], 'api', d['api_e2e']['shape_d']['value'])"
done
python -m pytest tests/test_multi_gpu.py -m gpu -q 2>&1 | tail -3
