#!/bin/bash
for cfg in "1 0" "2 0" "3 0" "4 0" "4 0.1"; do set -- $cfg
  echo "== chunks=$1 edge=$2"
  BENCH_E2E_CHUNKS=$1 BENCH_E2E_EDGE=$2 timeout 200 python bench.py --steps 200 --warmup 5 --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])"
done
