python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for c in 4 8 16; do CTR_HOST_CHUNKS=$c python scripts/e2e_bench.py 2>&1 | tail -1; done
E2E_PACKED=0 CTR_HOST_CHUNKS=8 python scripts/e2e_bench.py 2>&1 | tail -1
