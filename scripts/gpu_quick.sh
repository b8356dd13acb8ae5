for f in gym-ctr-reach_b200/variants/lib_*.so; do CTR_B200_LIB=$PWD/$f python scripts/kbench.py 2>&1 | tail -1; done
