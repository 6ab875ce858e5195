# cfg3 (2^27 vocab, K=64): record exchange against atomics exchange, same box, back to back.  usage: bash ... N
N=${1:-8}
for m in merge peer; do
  FM_AHEAD_PRIORITY=-1 timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --workload cfg3 --shard-mode $m 2>gpurun_out/cfg3_n${N}_$m.err | tail -1 > gpurun_out/cfg3_n${N}_$m.json
done
