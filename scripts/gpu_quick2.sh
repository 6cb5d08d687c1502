python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/op_timing.py 1000
