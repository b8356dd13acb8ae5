# steady-state step time against the number of leading refill blocks ($CTR_REFILL_BLOCKS; default = SM count)
cd "$(dirname "$0")/.."
out=gpurun_out/refill_sweep.txt; : > $out
for rep in 1 2; do for rb in 0 37 74 148 296 592; do
CTR_REFILL_BLOCKS=$rb python bench.py --steps 100 --warmup 10 --no-cpu --no-kernel-table --no-extra 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('refill_blocks %4d  value %.4g ms/step %.4f e2e %.4g reset_fraction %.4f' % ($rb, d['value'], d['ms_per_step'], d['e2e']['value'], d['reset_fraction']))" | tee -a $out
done; done
