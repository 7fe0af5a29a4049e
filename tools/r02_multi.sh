#!/bin/bash
# Round-2 multi-GPU session: step variants, parity on real peers, bench (+ reference arm), probes.
# usage: tools/r02_multi.sh N [tag]   (run on the GPU box from the repo root)
N=${1:-2}; TAG=${2:-r02_${N}gpu}
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { name=$1; shift; echo "== $name"; timeout ${T:-300} "$@" > $OUT/${TAG}_$name.log 2> $OUT/${TAG}_$name.err; echo "rc=$?"; grep -v Warning $OUT/${TAG}_$name.log | tail -n ${TAILN:-12}; grep -v -i "warn\|symm_mem\|\*\*\*\|OMP_NUM" $OUT/${TAG}_$name.err | tail -n 6; }
if [ -z "$SKIP_SWEEP" ]; then T=400 TAILN=60 run sweep $TR --master-port 29510 tools/step_sweep.py ${SWEEP:-headline small}; fi
if [ -z "$SKIP_VERIFY" ]; then T=400 TAILN=60 VERIFY_ONLY=${VERIFY_ONLY:-fused,switch,switch-inplace,host} run verify $TR --master-port 29511 tools/verify_sharded.py; fi
T=700 TAILN=2 run bench_default $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5
TAILN=2 run bench_ref python bench.py --impl reference --gpus $N --steps 5 --warmup 2
if [ -n "$WITH_PROBE" ]; then T=240 TAILN=40 PROBE_CTAS=296,592 run probe $TR --master-port 29515 tools/nvls_probe.py; fi
