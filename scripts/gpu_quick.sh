# quick check of a kernel change: GPU parity tests, then a short bench line (no CPU baseline)
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --steps 50 --warmup 5 --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo rc=$?
tail -c 800 gpurun_out/bench_quick.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_quick.json') if l.startswith('{')][-1])
print("value %.4g e2e %.4g ms/step %.4f k_ms %.4f frac %.3f peak %.2f"%(d['value'],d['e2e']['value'],d['ms_per_step'],d['roofline']['kernel_ms'],d['roofline']['frac'],d['roofline']['peak']))
PY
