lscpu | grep -i "numa\|^CPU(s)\|socket" ; nvidia-smi topo -m 2>&1 | head -20
for i in 0 1; do cat /sys/bus/pci/devices/$(nvidia-smi --query-gpu=pci.bus_id --format=csv,noheader -i $i | tr 'A-Z' 'a-z' | sed 's/^0000//')/numa_node 2>/dev/null; done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 2>&1 | tail -3
