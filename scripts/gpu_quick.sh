set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/kbench.py 2>&1 | tail -1
