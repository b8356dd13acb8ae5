for lib in "" s6 s8; do for c in 6 8 12 16; do
  if [ -z "$lib" ]; then L=""; else L="CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$lib.so"; fi
  echo -n "streams=${lib:-s4} "; env $L CTR_HOST_CHUNKS=$c python scripts/e2e_bench.py 2>&1 | tail -1
done; done
