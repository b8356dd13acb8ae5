# ctr_step_host with and without its kernels (scripts/build_variant.sh nokernel -DCTR_DEBUG_SKIP_STEP_KERNEL first): what the copies alone cost
for c in 1 4 8 16; do
  echo -n "with kernels   "; CTR_HOST_CHUNKS=$c python scripts/e2e_bench.py 2>&1 | tail -1
  echo -n "copies only    "; CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_nokernel.so CTR_HOST_CHUNKS=$c python scripts/e2e_bench.py 2>&1 | tail -1
done
