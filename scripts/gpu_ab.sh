# A/B of library variants with scripts/kbench.py (k_step alone, L2 flushed, and step + auto-reset), then the GPU parity tests
# on the in-tree library.  usage: gpu_ab.sh [variant names under gym-ctr-reach_b200/variants/lib_<name>.so ...]
cd "$(dirname "$0")/.."
out=gpurun_out/ab.txt; : > $out
for rep in 1 2; do
  python scripts/kbench.py 2>&1 | tail -1 | tee -a $out
  for v in "$@"; do
    CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$v.so python scripts/kbench.py 2>&1 | tail -1 | tee -a $out
  done
done
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee -a $out
