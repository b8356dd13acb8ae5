set -x
python -m pytest tests -m gpu -q 2>&1 | tail -6
python bench.py --steps 30 --warmup 5 --no-cpu --extras > gpurun_out/bench_her.json 2> gpurun_out/bench_her.err; echo rc=$?
tail -c 1200 gpurun_out/bench_her.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_her.json').read())
for k,v in d['extra'].items():
    if k.startswith('her') or k.startswith('cfg4'): print(k, json.dumps(v))
PY
