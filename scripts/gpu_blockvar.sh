for v in "" b256 b512 "" b256; do
  if [ -z "$v" ]; then python scripts/kbench.py 2>&1 | tail -1; else CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$v.so python scripts/kbench.py 2>&1 | tail -1; fi
done
