python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python scripts/kbench.py 2>&1 | tail -1
CTR_NO_SORT=1 python scripts/kbench.py 2>&1 | tail -1
