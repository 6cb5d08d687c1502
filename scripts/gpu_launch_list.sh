# ncu launch list of the bench command (one GPU): every kernel launch of a short bench run with its duration.
# Numbers printed under ncu are never bench values.
python bench.py --steps 2 --warmup 3 --seeds 0 --no-cpu-baseline > gpurun_out/plain_launches.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_r02.csv \
    python bench.py --steps 2 --warmup 3 --seeds 0 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log; wc -l gpurun_out/launches_r02.csv
