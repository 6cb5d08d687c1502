python -m pytest tests/test_rng_parity.py -x -q 2>&1 | grep -E "Error|assert|passed|failed" | head
ncu --set full --clock-control none --import-source on -k regex:mcmc_step -s 5 -c 1 -f -o gpurun_out/k8_op_5b python scripts/op_timing.py 300 > gpurun_out/k8_op_5b.log 2>&1
