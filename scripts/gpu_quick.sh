# quick check of a K8 change: per-operator timing, GPU tests, short bench (no CPU baseline / likelihood leg)
set -x
python scripts/op_timing.py 2000 > gpurun_out/op_timing.log 2>&1; cat gpurun_out/op_timing.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-likelihood-leg --no-variants > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_quick.json'));print({k:d[k] for k in ('value','ms_per_step','us_per_proposal_per_chain','ll_drift_rel','acceptance')});print(d['e2e'])"
