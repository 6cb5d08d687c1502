# Round-2 closing run (one gpurun call, 1 GPU): full GPU suite, bench line + reference arm with the final library, and
# one ncu --set full pass over the three blocked evaluation kernels at shape D (K1 lh_free_block, K1s lh_source_block,
# K4 count_block): the command's first 64 count_block launches are the one-chain recounts of sb_set_state, so the
# capture window (-s 263 -c 3 over the matched kernels: 2 x 64 one-chain recounts + 3 x 45 launches come first) is the closing K1, K1s, K4 launch of scripts/k1_shape_d.py.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_final.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r02_ref.json 2>> gpurun_out/bench_r02.err
python scripts/k1_shape_d.py > gpurun_out/kd_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"lh_free_block|lh_source_block|count_block" -s 263 -c 3 -f -o gpurun_out/kd_blocked \
    python scripts/k1_shape_d.py > gpurun_out/kd_ncu_blocked.log 2>&1
cat gpurun_out/kd_plain.log
