# ncu --set full captures of the secondary kernels (each after its driver exited 0 without ncu)
set -x
python scripts/herbench.py > gpurun_out/her_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_her_sample -s 3 -c 1 -f -o gpurun_out/prof_her_sample python scripts/herbench.py > gpurun_out/ncu_her1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_her_store -s 20 -c 1 -f -o gpurun_out/prof_her_store python scripts/herbench.py > gpurun_out/ncu_her2.log 2>&1
python scripts/shapebench.py > gpurun_out/shape_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_fk_shape -s 1 -c 1 -f -o gpurun_out/prof_fk_shape python scripts/shapebench.py > gpurun_out/ncu_shape.log 2>&1
tail -2 gpurun_out/ncu_her1.log gpurun_out/ncu_her2.log gpurun_out/ncu_shape.log
ls -la gpurun_out/*.ncu-rep
