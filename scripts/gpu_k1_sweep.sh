for w in ${K1_SWEEP:-888 1036 2072}; do echo "== SB_K1_CTAS=$w"; SB_K1_CTAS=$w python scripts/k1_only.py 30 2>&1 | cut -c1-140; done
