timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 500 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo rc=$?; tail -5 gpurun_out/bench_q.err
python -c "
import json; d=json.load(open('gpurun_out/bench_q.json')); print(d['roofline']); [print(x) for x in d['likelihood_sweep']]; print(d['value'], d['kernel_ms_per_step'], d['e2e'])"
