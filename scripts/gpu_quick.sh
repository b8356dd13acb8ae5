python -m pytest tests -m gpu -x -q 2>&1 | tail -15
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/libctr_b200.so python scripts/kbench.py 2>&1 | tail -1
for f in gym-ctr-reach_b200/variants/lib_*.so; do [ -f $f ] && CTR_B200_LIB=$PWD/$f python scripts/kbench.py 2>&1 | tail -1; done
KB_PREC=f32 python scripts/kbench.py 2>&1 | tail -1
KB_INTEG=rk4 python scripts/kbench.py 2>&1 | tail -1
