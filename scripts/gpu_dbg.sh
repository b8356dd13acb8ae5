echo "== A: test_her alone"; python -m pytest tests/test_her.py -m gpu -q 2>&1 | tail -4
echo "== B: error_paths then her"; python -m pytest "tests/test_gpu_parity.py::test_error_paths" tests/test_her.py -m gpu -q 2>&1 | tail -4
echo "== C: host_step test then her"; python -m pytest "tests/test_gpu_parity.py::test_host_step_entry_point_matches_device_path" tests/test_her.py -m gpu -q 2>&1 | tail -4
echo "== D: full, her first"; python -m pytest tests/test_her.py tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -4
