set -x
python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 4 -c 1 -f -o gpurun_out/prof_step python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
