# ncu --set full of single-operator launches of the step kernel (launch index = position in scripts/op_timing.py)
for idx in 3 11; do
  ncu --set full --clock-control none --import-source on -k regex:mcmc_step -s $idx -c 1 -f -o gpurun_out/k8_op_$idx \
      python scripts/op_timing.py 1000 > gpurun_out/k8_op_$idx.log 2>&1
done
ls -la gpurun_out/k8_op_*
