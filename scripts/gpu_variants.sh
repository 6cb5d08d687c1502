for v in A B C; do for t in 1 2; do echo "variant $v team $t"; SBAYES_B200_TEAM=$t SBAYES_B200_LIB=contact_zones_b200/csrc/variants/lib_$v.so python scripts/op_timing.py 1000 | head -5; done; done
