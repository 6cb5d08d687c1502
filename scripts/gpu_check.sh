# Parity + bench check (one gpurun call): GPU tests, full bench line, per-operator timing.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench.json
python scripts/op_timing.py 1000 > gpurun_out/op_timing.log 2>&1; tail -30 gpurun_out/op_timing.log
