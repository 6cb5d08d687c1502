for v in "$@"; do echo "== $v"; SBAYES_B200_LIB=variants/$v python scripts/k1_shape_d.py 2>&1 | tail -3; done
for c in 512 640 768 850 896; do echo "== SB_K1_CTAS_COMPACT=$c"; SB_K1_CTAS_COMPACT=$c python scripts/k1_shape_d.py 2>&1 | grep "^k1 "; done
