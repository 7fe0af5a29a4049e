#!/bin/bash
# Sharded-step variants at N GPUs: usage tools/bench_sharded.sh N [steps]
N=$1; K=${2:-30}
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps $K --warmup 5 "$@" 2>&1 | grep -E "^\{|Error" | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); c=d['config']; print(c['mode'], c['rendezvous'], 'mc' if c['multicast'] else 'uc', 'chunks', c['chunks'], ':', d['ms_per_step'], 'ms', d['value'], 'GB/s')
    else: print(l.strip()[:200])
"; }
shift; shift
run --mode fused --chunks 1
run --mode fused --chunks 4
run --mode fused --chunks 8
run --mode fused --chunks 16
run --mode fused --chunks 32
run --mode fused --chunks 64
run --mode streams --chunks 1
run --mode streams --chunks 2
run --mode nccl
