python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/op_timing.py 1000
SBAYES_B200_LIB=contact_zones_b200/csrc/libsbayes_b200_prof.so python scripts/op_phases.py 500
