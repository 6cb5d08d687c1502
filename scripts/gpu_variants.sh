echo "== 1024 x team 2"; SBAYES_B200_LIB=contact_zones_b200/csrc/variants/lib_t1024.so python scripts/op_timing.py 2000 2>&1 | head -12
