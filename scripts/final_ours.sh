#!/bin/bash
# this framework only, sustained: scripts/final_ours.sh N [steps] -> gpurun_out/final_b200dp_N.json
N="$1"; STEPS="${2:-1000}"
mkdir -p gpurun_out
run() { if [ "$N" = "1" ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) "$@"; fi; }
run bench.py --gpus $N --steps $STEPS --warmup 10 2> gpurun_out/final_b200dp_$N.err | grep '^{"metric' | tail -1 > gpurun_out/final_b200dp_$N.json
python -c "
import json; d=json.load(open('gpurun_out/final_b200dp_$N.json')); print($N, round(d['value']), round(d['e2e']['value']), round(d['ms_per_step'],4), d['clocks'], d['comm'])"
