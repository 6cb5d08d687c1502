for t in 1 2; do
SBAYES_B200_TEAM=$t python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-likelihood-leg > gpurun_out/bench_team$t.json 2> gpurun_out/bench_team$t.err; tail -2 gpurun_out/bench_team$t.err
python -c "
import json;d=json.load(open('gpurun_out/bench_team$t.json'));print('team $t', {k:d[k] for k in ('value','ms_per_step','us_per_proposal_per_chain','ll_drift_rel','acceptance')});print(d['e2e'])"
done
