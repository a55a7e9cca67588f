#!/bin/bash
# facts about the GPU box's host side (for the multi-GPU end-to-end numbers): sockets, NUMA nodes, allowed cpus, GPU nodes
echo "== nproc: $(nproc); allowed cpus: $(taskset -cp $$ 2>/dev/null | sed 's/.*: //')"
for n in /sys/devices/system/node/node*; do echo "== $(basename $n): cpus $(cat $n/cpulist); $(grep MemTotal $n/meminfo | awk '{print $4/1048576 " GB"}')"; done
nvidia-smi --query-gpu=index,pci.bus_id --format=csv,noheader | while IFS=, read idx bdf; do b=$(echo $bdf | tr 'A-Z' 'a-z' | sed 's/^ *0000//'); echo "== gpu $idx $bdf numa_node $(cat /sys/bus/pci/devices/$b/numa_node 2>/dev/null)"; done
nvidia-smi topo -m 2>/dev/null | head -14
lscpu | grep -E "Model name|Socket|NUMA|Thread|Core" 
