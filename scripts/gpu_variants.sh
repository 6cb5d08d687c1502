# per-operator timing of step-kernel build variants (contact_zones_b200/csrc/variants/lib_*.so), first lines only
for v in ${VARIANTS}; do
  echo "== $v"; SBAYES_B200_LIB=contact_zones_b200/csrc/variants/lib_$v.so python scripts/op_timing.py 2048 2>&1 | head -${LINES:-4}
done
