python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
SB_K1C_MIN_CHAINS=1 SB_K1_BLOCK_MAX_N=0 SB_FUZZ_SEEDS_K1=40 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_model_api.py -m gpu -x -q 2>&1 | tail -3
echo "== default (K1c up to 8192 sites)"; python scripts/k1_only.py 30 500,1000,2000,5000,10000,20000 2>&1 | python -c "
import sys,ast
for l in sys.stdin:
    if l.startswith('{'):
        r=ast.literal_eval(l); print(r['N'], round(r['us_mean'],1))
"
echo "== K1c everywhere"; SB_K1C_MAX_N=1000000 python scripts/k1_only.py 30 10000,20000 2>&1 | python -c "
import sys,ast
for l in sys.stdin:
    if l.startswith('{'):
        r=ast.literal_eval(l); print(r['N'], round(r['us_mean'],1))
"
