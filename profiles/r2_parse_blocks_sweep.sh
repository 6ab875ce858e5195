# grid cap of the run-ahead parse, again with the round-2 parser (profiles/r1_parse_blocks_sweep.txt was the slower one)
for dist in uniform zipf; do for pb in 148 222 296 444; do
  timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --dist $dist --parse-blocks $pb 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$dist', $pb, round(d['value']/1e6,1), round(d['ms_per_step'],4), r['stages_pipelined_ms'])"
done; done
