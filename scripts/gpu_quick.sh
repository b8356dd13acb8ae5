set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python bench.py --steps 30 --warmup 5 --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo bench rc=$?
tail -c 2500 gpurun_out/bench_quick.json; tail -5 gpurun_out/bench_quick.err
