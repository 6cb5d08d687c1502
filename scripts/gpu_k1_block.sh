python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_model_api.py tests/test_results_golden.py -m gpu -x -q 2>&1 | tail -2
python scripts/k1_only.py 30 500,1000,2000,5000 2>&1 | python -c "
import sys,ast
for l in sys.stdin:
    if l.startswith('{'):
        r=ast.literal_eval(l); print(r['N'], round(r['us_mean'],1))
"
python scripts/k1_shape_d.py 2>&1 | grep "^k1"
