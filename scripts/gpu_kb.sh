# kbench A/B only: in-tree library + the named variants, REPS rounds (default 2).  usage: gpu_kb.sh name ...
cd "$(dirname "$0")/.."
out=gpurun_out/kb.txt; : > $out
for rep in $(seq ${REPS:-2}); do
  python scripts/kbench.py 2>&1 | tail -1 | tee -a $out
  for v in "$@"; do
    CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$v.so python scripts/kbench.py 2>&1 | tail -1 | tee -a $out
  done
done
