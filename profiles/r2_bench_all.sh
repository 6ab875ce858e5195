#!/bin/bash
# round-2 single-GPU measurements: bench.py per path and id distribution, the reference arm, the bulk-copy probe
mkdir -p gpurun_out
for d in uniform zipf; do for p in claim sorted; do
  timeout 300 python bench.py --steps 40 --warmup 5 --dist $d --path $p $([ "$p$d" != claimuniform ] && echo --no-cpu-baseline) \
    > gpurun_out/bench_r2_n1_${p}_${d}.json 2> gpurun_out/bench_r2_n1_${p}_${d}.err
done; done
timeout 300 python bench.py --impl reference --steps 8 --warmup 1 > gpurun_out/bench_r2_reference_arm.json 2>&1
timeout 120 ./profiles/bulk_rmw_probe.bin > gpurun_out/r2_bulk_rmw_probe.jsonl 2>&1
for f in gpurun_out/bench_r2_n1_*.json; do python - "$f" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read()); r = d["roofline"]
print(sys.argv[1], "step %.4f e2e %.4f U %d" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["unique_per_step"]),
      {k: (v["ms"], v["frac"], v["frac_as_implemented"]) for k, v in r["stages"].items()})
PY
done
