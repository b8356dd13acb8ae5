# extras incl. the config-5 curriculum and config-3 tol 0.001 lines, then the torchrun path at N = 2 with the current bench.py
python bench.py --steps 30 --warmup 5 --no-cpu --no-kernel-table --extras > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo rc=$?
tail -c 600 gpurun_out/bench_x.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_x.json').read())
for k,v in d['extra'].items():
    if k.startswith('cfg'): print(k, v.get('value', v.get('us_per_call')))
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/v2b.out 2> gpurun_out/v2b.err; echo rc=$?
echo "stdout lines: $(wc -l < gpurun_out/v2b.out)"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/v2b.out').read())
print("N=%d value %.4g e2e %.4g keys %s"%(d['n_gpus'],d['value'],d['e2e']['value'],sorted(d.keys())))
PY
