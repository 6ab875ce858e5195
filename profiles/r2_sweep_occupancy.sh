#!/bin/bash
# experiment: cap the table kernels' occupancy (unused dynamic shared memory) so that the run-ahead parse keeps a slot
for cfg in "0 0 296" "75 0 296" "0 45 296" "75 45 296" "75 45 148" "75 56 148" "113 45 296" "75 45 592"; do
  set -- $cfg
  echo "== score_cap_kb=$1 claim_cap_kb=$2 parse_blocks=$3"
  FM_SCORE_SMEM_CAP_KB=$1 FM_CLAIM_SMEM_CAP_KB=$2 timeout 120 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --parse-blocks $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('step %.4f e2e %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step']), {k:v['ms'] for k,v in r['stages'].items()}, r['stages_pipelined_ms'])"
done
