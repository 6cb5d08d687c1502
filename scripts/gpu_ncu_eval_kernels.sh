# ncu --set full of the blocked full-evaluation kernels (K4 recount, K1 with source) as launched by bench.py at shape D
ncu --set full --clock-control none --import-source on -k regex:count_block -s 2 -c 1 -f -o gpurun_out/k4_count_block \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-likelihood-leg --no-variants --e2e-steps 1 > gpurun_out/k4_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lh_source_block -s 1 -c 1 -f -o gpurun_out/k1s_source_block \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-likelihood-leg --no-variants --e2e-steps 1 > gpurun_out/k1s_ncu.log 2>&1
ls -la gpurun_out/k4_count_block.ncu-rep gpurun_out/k1s_source_block.ncu-rep
