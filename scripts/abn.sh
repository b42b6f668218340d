#!/bin/bash
# multi-GPU A/B: scripts/abn.sh N "<label>" <bench args...>
n="$1"; label="$2"; shift; shift
out=$(python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 100 --no-e2e "$@" 2>&1 | grep '^{"metric"' | tail -1)
echo "$label: $(echo "$out" | python -c "import sys,json
try:
    d=json.loads(sys.stdin.read()); print('%.4f ms/step  %.0f img/s  %s' % (d['ms_per_step'], d['value'], d.get('comm')))
except Exception as e:
    print('FAILED', e)")"
