# A/B of the lane-exchange mechanism: default library (CTR_XCH_SMEM=1) against variants/lib_shfl.so (-DCTR_XCH_SMEM=0), then the
# whole GPU suite on the default library.
for rep in 1 2 3; do
python scripts/kbench.py 2>&1 | tail -1
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_shfl.so python scripts/kbench.py 2>&1 | tail -1
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
