# weak-scaling bench at N GPUs of one box (launched the way the driver does): bash scripts/gpu_scale.sh "2 4 8"
for n in ${1:-8}; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520 + n)) \
      bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "N=$n rc=$? lines=$(wc -l < gpurun_out/bench_n$n.json)"
  python -c "
import json;d=json.load(open('gpurun_out/bench_n$n.json'));print({k:d[k] for k in ('value','n_gpus','ms_per_step','kernel_ms_per_step','scaling')});print(d['e2e'])"
done
