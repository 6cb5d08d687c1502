python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_model_api.py -m gpu -x -q 2>&1 | tail -2
for t in 256 192 160 128 96; do echo "== threads $t"; SB_K1B_THREADS=$t SB_K1S_THREADS=$t python scripts/k1_shape_d.py 2>&1 | grep "^k1"; done
