python scripts/pcie_probe.py 2>&1 | grep -v Warning | tail -4
for n in 2 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 scripts/pcie_probe.py 2>&1 | grep -E "world|rank|aggregate"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29522 scripts/pcie_probe.py --bind 2>&1 | grep -E "world|rank|aggregate"
done
nvidia-smi topo -m 2>&1 | head -14
lscpu | grep -E "Model name|Socket|NUMA node|^CPU\(s\)"
