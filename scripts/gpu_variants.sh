echo "== base 512 x team 2"; python scripts/op_timing.py 1000 2>&1 | head -6
echo "== 256 x team 4 (2 CTAs/SM)"; SBAYES_B200_LIB=contact_zones_b200/csrc/variants/lib_t256.so SBAYES_B200_TEAM=4 python scripts/op_timing.py 1000 2>&1 | head -9
echo "== 256 x team 2"; SBAYES_B200_LIB=contact_zones_b200/csrc/variants/lib_t256.so SBAYES_B200_TEAM=2 python scripts/op_timing.py 1000 2>&1 | head -6
