# Round-2 evidence pass of the current kernels: GPU parity tests, smoke, bench line + reference arm, ncu launch list and full-set
# captures of k_step and k_her_sample (each ncu pass only after the same command exited 0 without ncu).
T=${TAG:-r2m}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 50 --warmup 5 > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo bench rc=$?
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref_$T.json 2>/dev/null; echo ref rc=$?
Q="--steps 3 --warmup 3 --no-cpu --no-kernel-table --no-extra"
python bench.py $Q > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_$T.csv python bench.py $Q > gpurun_out/ncu1.log 2>&1
python bench.py $Q > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 12 -c 1 -f -o gpurun_out/prof_step_$T python bench.py $Q > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
python scripts/herbench.py > gpurun_out/her_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_her_sample -s 3 -c 1 -f -o gpurun_out/prof_her_sample_$T python scripts/herbench.py > gpurun_out/ncu_her1.log 2>&1
tail -1 gpurun_out/ncu_her1.log
ls -la gpurun_out/*.ncu-rep
