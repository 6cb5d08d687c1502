set -e
python scripts/k1_only.py 20 > gpurun_out/k1_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lh_table -s 60 -c 2 -f -o gpurun_out/k1_prof python scripts/k1_only.py 20 > gpurun_out/k1_ncu.log 2>&1
cat gpurun_out/k1_plain.log
tail -3 gpurun_out/k1_ncu.log
