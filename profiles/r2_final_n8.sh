# final code at 8 GPUs: cfg2 and cfg3, default (atomics) exchange
for w in cfg2 cfg3; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 40 --warmup 5 --workload $w 2>gpurun_out/n8_final_$w.err | tail -1 > gpurun_out/n8_final_$w.json
done
