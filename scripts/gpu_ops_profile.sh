# ncu --set full of single-operator launches of the step kernel.  Launch index = 2 * position in scripts/op_timing.py's
# mix list + 1 (each mix launches a 20-step warm-up and the timed run): 3 gibbs_mix, 11 p_zones, 17 grow, 23 sources
for idx in ${OPS_IDX:-23 17}; do
  ncu --set full --clock-control none --import-source on -k regex:mcmc_step -s $idx -c 1 -f -o gpurun_out/k8_op_$idx \
      python scripts/op_timing.py 1000 > gpurun_out/k8_op_$idx.log 2>&1
done
ls -la gpurun_out/k8_op_*
