python bench.py --steps 30 --warmup 5 --no-cpu > gpurun_out/bench_ktab.json 2> gpurun_out/bench_ktab.err; echo rc=$?
tail -c 1500 gpurun_out/bench_ktab.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_ktab.json').read())
print("k_step frac %.3f  %.3f ms"%(d['roofline']['frac'], d['roofline']['kernel_ms']))
for k,v in d['kernels'].items(): print("%-22s %-4s %9.1f us  %8.2f / %8.2f %-8s frac %.3f"%(k,v['bound'],v['us'],v['achieved'],v['peak'],v['unit'],v['frac']))
PY
