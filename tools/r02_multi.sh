#!/bin/bash
# Round-2 multi-GPU session: parity on real peers, bench variants, link probes.
# usage: tools/r02_multi.sh N [tag]   (run on the GPU box from the repo root)
N=${1:-2}; TAG=${2:-r02_${N}gpu}
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; echo "== $name"; timeout ${T:-300} "$@" > $OUT/${TAG}_$name.log 2> $OUT/${TAG}_$name.err; echo "rc=$?"; tail -n ${TAILN:-12} $OUT/${TAG}_$name.log; tail -c 1500 $OUT/${TAG}_$name.err | grep -v "^$" | tail -n 8; }
T=400 TAILN=40 run verify $TR --master-port 29511 tools/verify_sharded.py
T=500 TAILN=3 run bench_default $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5
TAILN=3 run bench_switch8 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --mode switch --chunks 8 --no-configs --no-e2e
TAILN=3 run bench_switch16 $TR --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --mode switch --chunks 16 --no-configs --no-e2e
TAILN=3 run bench_ref python bench.py --impl reference --gpus $N --steps 5 --warmup 2
T=240 TAILN=40 run probe $TR --master-port 29515 tools/nvls_probe.py
