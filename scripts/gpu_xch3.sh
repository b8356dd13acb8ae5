# A/B: default library against a variant ($V), then the GPU suite on the variant (minus the in-tree-path assertion)
V=${V:-trim}
for rep in 1 2 3; do
python scripts/kbench.py 2>&1 | tail -1
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$V.so python scripts/kbench.py 2>&1 | tail -1
done
CTR_B200_LIB=$PWD/gym-ctr-reach_b200/variants/lib_$V.so timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_parity.py::test_library_loaded_is_in_tree 2>&1 | tail -3
