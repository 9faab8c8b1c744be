#!/bin/bash
# Tuning aid: cost of the pieces of the peer-memory exchange at N GPUs (per-step time with the exchange disabled, with
# only its barrier, and complete).
N=${1:-4}
show() { python -c "
import json,sys
j=json.loads(open('$1').read()); print('$2', 'value %.1f G/s'%(j['value']/1e9), 'ms %.3f'%j['ms_per_step'], 'kernel_ms %.3f'%j['roofline']['kernel_ms'])"; }
for dbg in setup-only barriers-only ""; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --exchange peer --exchange-debug "$dbg" 2>/dev/null | grep metric > /tmp/nn.json; show /tmp/nn.json "N=$N peer dbg=$dbg"
done
