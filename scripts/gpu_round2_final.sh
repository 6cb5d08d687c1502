# Round-2 closing run (one gpurun call, 1 GPU): full GPU suite, bench line + reference arm with the final library, and
# the ncu --set full capture of K4 (count_block) on a 64-chain launch at shape D: the first 64 launches of the command
# are the one-chain recounts of sb_set_state, then 5 warm-up launches, so launch 70+ is a timed 64-chain one.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_final.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r02_ref.json 2>> gpurun_out/bench_r02.err
python scripts/k1_shape_d.py > gpurun_out/kd_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:count_block -s 72 -c 1 -f -o gpurun_out/kd_count_block \
    python scripts/k1_shape_d.py > gpurun_out/kd_ncu_count_block.log 2>&1
cat gpurun_out/kd_plain.log
