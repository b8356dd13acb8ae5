# ctr_step_host with / without descending stream priorities, over chunk counts
for c in 8 12 16; do for p in 0 1; do
echo -n "prio=$p "; CTR_HOST_PRIO=$p CTR_HOST_CHUNKS=$c python scripts/e2e_bench.py 2>&1 | tail -1
done; done
echo -n "prio=1 nograph "; CTR_HOST_GRAPH=0 CTR_HOST_PRIO=1 CTR_HOST_CHUNKS=8 python scripts/e2e_bench.py 2>&1 | tail -1
python -m pytest tests -m gpu -x -q -k "host or destroy" 2>&1 | tail -2
