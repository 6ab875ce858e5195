# record exchange against the atomics exchange at N=8, same box, back to back
for m in merge peer; do
  FM_AHEAD_PRIORITY=-1 timeout 250 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 30 --warmup 5 --shard-mode $m 2>gpurun_out/n8_$m.err | tail -1 > gpurun_out/n8_$m.json
done
