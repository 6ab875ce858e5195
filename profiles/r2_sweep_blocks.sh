for u in 2 3 4 6; do for s in 8 4; do
  echo "== update_bpsm=$u score_bpsm=$s"
  FM_UPDATE_BLOCKS_PER_SM=$u FM_SCORE_BLOCKS_PER_SM=$s timeout 120 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('step %.4f e2e %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step']), {k:v['ms'] for k,v in r['stages'].items()}, r['stages_pipelined_ms'])"
done; done
