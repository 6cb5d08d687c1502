# per-GPU load of the strong-scaling line at N = 8 / 4 / 2 (8 / 16 / 32 chains on one GPU) for each team size
for nc in 8 16 32; do for t in 2 4 8 16; do
  echo "== chains $nc team $t"
  SBAYES_B200_TEAM=$t python bench.py --steps 3 --warmup 3 --chains-total $nc --seeds 0 --no-cpu-baseline --no-likelihood-leg --no-variants --no-api-leg --e2e-steps 2 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print({k:d[k] for k in ('value','kernel_ms_per_step','us_per_proposal_per_chain')})"
done; done
