python scripts/k1_only.py 20 > gpurun_out/k1_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lh_table -s 10 -c 1 -f -o gpurun_out/k1_small_prof python scripts/k1_only.py 20 > gpurun_out/k1_ncu.log 2>&1
tail -2 gpurun_out/k1_ncu.log
