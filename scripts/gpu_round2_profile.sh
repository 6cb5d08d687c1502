# Round-2 evidence (one gpurun call): bench line, reference arm, ncu launch list of the bench command, ncu --set full of
# the step kernel at 64 and at 8 chains per GPU, of K1 at three sweep sizes and of K1 / K1s / K4 at shape D.
# Numbers printed under ncu are never bench values.
set -x
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r02_ref.json 2>> gpurun_out/bench_r02.err
python bench.py --steps 2 --warmup 3 --seeds 0 --no-cpu-baseline > gpurun_out/plain_launches.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_r02.csv \
    python bench.py --steps 2 --warmup 3 --seeds 0 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python scripts/k8_only.py 64 > gpurun_out/k8_plain64.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mcmc_step -s 4 -c 1 -f -o gpurun_out/k8_step_64 \
    python scripts/k8_only.py 64 > gpurun_out/k8_ncu64.log 2>&1
python scripts/k8_only.py 8 > gpurun_out/k8_plain8.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mcmc_step -s 4 -c 1 -f -o gpurun_out/k8_step_8 \
    python scripts/k8_only.py 8 > gpurun_out/k8_ncu8.log 2>&1
python scripts/k1_only.py 20 500,2000,20000 > gpurun_out/k1_plain.log 2>&1 && \
for n in 500 2000 20000; do
ncu --set full --clock-control none --import-source on -k regex:lh_table -s 12 -c 1 -f -o gpurun_out/k1_N$n \
    python scripts/k1_only.py 20 $n > gpurun_out/k1_ncu_$n.log 2>&1
done
python scripts/k1_shape_d.py > gpurun_out/kd_plain.log 2>&1 && \
for k in lh_table lh_source_block count_block; do
ncu --set full --clock-control none --import-source on -k regex:$k -s 12 -c 1 -f -o gpurun_out/kd_$k \
    python scripts/k1_shape_d.py > gpurun_out/kd_ncu_$k.log 2>&1
done
ls -la gpurun_out | tail -20
