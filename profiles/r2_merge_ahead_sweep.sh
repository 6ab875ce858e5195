# record exchange at N=2: how the run-ahead parse + sort should be scheduled (stream priority, grid cap)
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 30 --warmup 5 --shard-mode merge 2>gpurun_out/ahead_err.log | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['stage_ms_rank0'])"; }
for pr in 0 -1; do for bl in 296 592 0; do echo "priority=$pr blocks=$bl"; FM_AHEAD_PRIORITY=$pr FM_AHEAD_BLOCKS=$bl run; done; done
