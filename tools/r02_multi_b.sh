#!/bin/bash
# Round-2 multi-GPU session B: the in-switch step after the lane-contiguous fix.
N=${1:-2}; TAG=${2:-r02b_${N}gpu}
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; echo "== $name"; timeout ${T:-300} "$@" > $OUT/${TAG}_$name.log 2> $OUT/${TAG}_$name.err; echo "rc=$?"; grep -v Warning $OUT/${TAG}_$name.log | tail -n ${TAILN:-12}; grep -v -i "warn\|symm_mem\|\*\*\*\|OMP_NUM" $OUT/${TAG}_$name.err | tail -n 6; }
T=300 TAILN=20 VERIFY_ONLY=switch run verify $TR --master-port 29511 tools/verify_sharded.py
T=400 TAILN=60 run sweep $TR --master-port 29510 tools/step_sweep.py headline small
T=200 TAILN=40 PROBE_MODES=7,8 PROBE_CTAS=37,74,148,296,592 run probe $TR --master-port 29515 tools/nvls_probe.py
if [ -n "$WITH_BENCH" ]; then
T=600 TAILN=2 run bench_default $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --no-big
fi
