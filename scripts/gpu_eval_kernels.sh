# parity of the evaluation kernels + their timing at shape D (one GPU)
python -m pytest tests/test_gpu_parity.py tests/test_sampler_invariants.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5
python scripts/k1_shape_d.py 2>&1 | tail -4
