# A/B of the steady-state step (`value` of bench.py: auto-reset at its steady load, L2 flushed) for library variants.
# usage: gpu_value_ab.sh name ...   (variants under gym-ctr-reach_b200/variants/lib_<name>.so; the in-tree library first)
cd "$(dirname "$0")/.."
out=gpurun_out/value_ab.txt; : > $out
run() { python bench.py --steps 100 --warmup 10 --no-cpu --no-kernel-table --no-extra 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('%-24s value %.4g ms/step %.4f e2e %.4g reset_fraction %.4f' % ('$1', d['value'], d['ms_per_step'], d['e2e']['value'], d['reset_fraction']))" | tee -a $out; }
for rep in 1 2; do
  run in-tree
  for v in "$@"; do CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$v.so run $v; done
done
