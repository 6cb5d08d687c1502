python -m pytest tests/test_step_features.py tests/test_sampler_invariants.py tests/test_sampler_parity.py -m gpu -x -q 2>&1 | tail -3
python scripts/op_timing_alter.py 2000 2>&1 | tail -9
python scripts/k8_only.py 64 2>&1 | tail -3
