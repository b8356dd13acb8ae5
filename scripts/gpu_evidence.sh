# Round-1 evidence pass of the current kernels: GPU parity tests, smoke, bench line (+extras, +reference arm),
# ncu launch list and one full-set capture of k_step (each ncu pass only after the same command exited 0 without ncu).
set -x
T=${TAG:-r1e}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --steps 100 --warmup 5 --extras > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo bench rc=$?
tail -c 1500 gpurun_out/bench_$T.err
python bench.py --impl reference --steps 100 --warmup 2 > gpurun_out/bench_ref_$T.json 2> gpurun_out/bench_ref_$T.err; echo ref rc=$?
cat gpurun_out/bench_ref_$T.json
python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$T.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu1.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 4 -c 1 -f -o gpurun_out/prof_step_$T python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
cat gpurun_out/bench_$T.json
