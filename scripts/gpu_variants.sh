# per-operator timing of step-kernel build variants (contact_zones_b200/csrc/variants/lib_*.so)
for v in t512_p0 t768_p1 t1024_p0 t1024_p1; do
  echo "== $v"; SBAYES_B200_LIB=contact_zones_b200/csrc/variants/lib_$v.so python scripts/op_timing.py 2000 2>&1 | head -11
done
