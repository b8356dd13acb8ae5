# ctr_step_host: done / success bytes per chunk (0) or in one copy after the last chunk (1, default)
for rep in 1 2; do for f in 0 1; do
echo -n "flags_late=$f "; CTR_HOST_FLAGS_LATE=$f python scripts/e2e_bench.py 2>&1 | tail -1
echo -n "flags_late=$f unpacked "; E2E_PACKED=0 CTR_HOST_FLAGS_LATE=$f python scripts/e2e_bench.py 2>&1 | tail -1
done; done
python -m pytest tests -m gpu -x -q -k "host or destroy or facade" 2>&1 | tail -2
