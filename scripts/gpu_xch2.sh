# A/B: default library against variants (pair = 128-bit mailbox rows; pair_desc = + heaviest-first window order)
for rep in 1 2 3; do
python scripts/kbench.py 2>&1 | tail -1
for v in pair pair_desc; do
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$v.so python scripts/kbench.py 2>&1 | tail -1
done
done
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_pair_desc.so timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
