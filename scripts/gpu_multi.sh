# Multi-GPU evidence: bench.py under torchrun at N = 2..$NG (one rank per GPU, NCCL only for the stats all-reduce)
set -x
NG=${NG:-2}
T=${TAG:-r1f}
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv
python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu > gpurun_out/bench_${T}_n1.json 2> gpurun_out/bench_${T}_n1.err; echo rc=$?
for n in 2 4 8; do
[ $n -le $NG ] || continue
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 50 --warmup 5 > gpurun_out/bench_${T}_n$n.json 2> gpurun_out/bench_${T}_n$n.err; echo rc=$?
tail -c 600 gpurun_out/bench_${T}_n$n.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $n --steps 20 --warmup 1 > gpurun_out/bench_ref_${T}_n$n.json 2>> gpurun_out/bench_${T}_n$n.err; echo rc=$?
done
cat gpurun_out/bench_${T}_n*.json | cut -c1-700
