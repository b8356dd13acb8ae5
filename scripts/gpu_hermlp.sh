# k_her_sample: records loaded per batch (CTR_HER_MLP) 1 (HEAD library) / 4 / 8 / 16, then the HER tests on the 8 variant
for rep in 1 2; do
python scripts/herbench.py 2>&1 | tail -1
for v in mlp4 mlp8 mlp16; do
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$v.so python scripts/herbench.py 2>&1 | tail -1
done
done
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_mlp8.so timeout 600 python -m pytest tests/test_her.py -m gpu -x -q 2>&1 | tail -3
