# Round-1 evidence run (one gpurun call): full bench, ncu launch list of the same command, ncu --set full of K1 at
# N=20000 and of the step kernel K8.  Numbers printed under ncu are never bench values.
set -x
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err || exit 1
tail -c 600 gpurun_out/bench_r01.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r01_ref.json 2>> gpurun_out/bench_r01.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r01.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python scripts/k1_only.py 20 20000 > gpurun_out/k1_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lh_table -s 12 -c 1 -f -o gpurun_out/k1_N20000 \
    python scripts/k1_only.py 20 20000 > gpurun_out/k1_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mcmc_step -s 2 -c 1 -f -o gpurun_out/k8_step \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-likelihood-leg --e2e-steps 1 > gpurun_out/k8_ncu.log 2>&1
ls -la gpurun_out | tail
