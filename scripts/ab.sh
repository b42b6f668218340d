#!/bin/bash
# A/B helper: scripts/ab.sh "<label>" <bench args...>  -> prints label and ms/step (device-timed, no e2e arm)
label="$1"; shift
out=$(python bench.py --steps 100 --no-e2e "$@" 2>&1 | tail -1)
ms=$(echo "$out" | python -c "import sys,json
try:
    d=json.loads(sys.stdin.read()); print('%.4f ms/step  %s' % (d['ms_per_step'], d.get('comm')))
except Exception as e:
    print('FAILED', e)")
echo "$label: $ms"
[ "$ms" = "${ms#FAILED}" ] || echo "$out" | tail -c 600
