python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/op_timing.py 1000
for idx in 3 5 15; do
  ncu --set full --clock-control none --import-source on -k regex:mcmc_step -s $idx -c 1 -f -o gpurun_out/k8_op_$idx \
      python scripts/op_timing.py 300 > gpurun_out/k8_op_$idx.log 2>&1
done
ls -la gpurun_out/k8_op_*
