python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; tail -3 gpurun_out/bench_q.err
python -c "
import json;d=json.load(open('gpurun_out/bench_q.json'));print({k:d[k] for k in ('value','ms_per_step','us_per_proposal_per_chain','ll_drift_rel','acceptance')});print(d['e2e']);print(d['roofline'])"
