V=${V:-estrin}
for rep in 1 2; do
python scripts/kbench.py 2>&1 | tail -1
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$V.so python scripts/kbench.py 2>&1 | tail -1
done
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$V.so python -m pytest tests/test_gpu_parity.py -m gpu -q -k "fk or rollout or step or reset" 2>&1 | tail -3
