# after a revert / bench change: GPU tests green again, stdout of the torchrun bench is exactly one JSON line, e2e at N=2
set -x
python -m pytest tests -m gpu -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/v2.out 2> gpurun_out/v2.err; echo rc=$?
echo "stdout lines: $(wc -l < gpurun_out/v2.out)"; head -c 60 gpurun_out/v2.out; echo
python - <<'PY'
import json
d=json.loads(open('gpurun_out/v2.out').read())
print("N=%d value %.4g e2e %.4g host_cpus %s k_ms %.4f frac %.3f"%(d['n_gpus'],d['value'],d['e2e']['value'],d['e2e'].get('host_cpus_rank0'),d['roofline']['kernel_ms'],d['roofline']['frac']))
PY
