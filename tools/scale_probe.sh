#!/bin/bash
# Tuning aid: per-GPU kernel time of bench.py at N=1 and N=$1 with both exchange forms (needs $1 GPUs).
N=${1:-4}
show() { python -c "
import json,sys
j=json.loads(open('$1').read()); print('$2', 'value %.1f G/s'%(j['value']/1e9), 'ms %.3f'%j['ms_per_step'], 'kernel_ms %.3f'%j['roofline']['kernel_ms'], j['clocks'])"; }
nvidia-smi --query-gpu=index,clocks.mem,clocks.max.mem,power.limit --format=csv | head -3
python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-mesh | grep metric > /tmp/n1.json; show /tmp/n1.json "N=1"
CUDA_VISIBLE_DEVICES=$((N-1)) python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-mesh | grep metric > /tmp/n1b.json; show /tmp/n1b.json "N=1 on last GPU"
for ex in nccl peer; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --exchange $ex 2>/dev/null | grep metric > /tmp/nn.json; show /tmp/nn.json "N=$N $ex"
done
# N independent single-GPU benches at the same time (no exchange at all)
for i in $(seq 0 $((N-1))); do CUDA_VISIBLE_DEVICES=$i python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-mesh | grep metric > /tmp/ind$i.json & done; wait
for i in $(seq 0 $((N-1))); do show /tmp/ind$i.json "independent $i"; done
