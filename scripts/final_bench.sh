#!/bin/bash
# All three arms at N GPUs, sustained (>= 5 s timed) for this framework: scripts/final_bench.sh N [steps]
# Writes gpurun_out/final_{ref,torch,b200dp}_N.json
N="$1"; STEPS="${2:-1000}"
mkdir -p gpurun_out
run() { if [ "$N" = "1" ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) "$@"; fi; }
run bench.py --impl reference --gpus $N --steps 20 --warmup 3 2> gpurun_out/final_ref_$N.err | grep '^{' | tail -1 > gpurun_out/final_ref_$N.json
STEPS=60 WARMUP=10 run baseline/torch_ddp_baseline.py 2> gpurun_out/final_torch_$N.err | grep '^{' | tail -1 > gpurun_out/final_torch_$N.json
run bench.py --gpus $N --steps $STEPS --warmup 10 2> gpurun_out/final_b200dp_$N.err | grep '^{"metric' | tail -1 > gpurun_out/final_b200dp_$N.json
for a in ref torch b200dp; do echo "== $a N=$N"; cut -c1-420 gpurun_out/final_${a}_$N.json; done
