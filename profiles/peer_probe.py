"""2+ GPU probe (torchrun): time fm_peer_score_scatter scoring-only vs training (remote reads vs remote atomics)."""
import os, sys, json
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_data
from bench import CFG
from fast_tffm_b200 import _lib
from fast_tffm_b200._lib import check, ptr, stream
from fast_tffm_b200.fm_model import Model
from fast_tffm_b200.fm_ops import PackedLines
from fast_tffm_b200.sharded import PeerShardedModel, _at

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
b, k = CFG["batch"], CFG["factor_num"]
model = Model(vocabulary_size=CFG["vocab"], factor_num=k, batch_size=b * world, hash_feature_id=True, loss_type="logistic")
m = PeerShardedModel(model, device=dev, max_nnz_per_line=39).build(seed=0)
text, off, _ = bench_data.criteo_batch(b, 1000 * (rank + 1))
batch = PackedLines(torch.from_numpy(bench_data.pad16(text)).to(dev), torch.from_numpy(off).to(dev))
packed, bufs = m._parse(batch)
lib, st = m.lib, stream()

def run(train):
    labels = ptr(bufs["labels"]) if train else None
    check(lib.fm_peer_score_scatter(ptr(bufs["ids"]), ptr(bufs["vals"]), ptr(bufs["poses"]), labels, None, b, k, world,
          m._table_bases, m.ld, m._gtab_bases if train else None, m.gld, m._touched_bases if train else None, 0.0, 0.0, 0, b * world,
          ptr(bufs["pred"]), ptr(bufs["scal"]), _at(bufs["scal"], 1), None, ptr(bufs["ws_fused"]), bufs["ws_fused"].numel(), st), "x")

res = {}
for name, train in (("score_only", False), ("score_scatter", True)):
    for _ in range(3):
        run(train)
    dist.barrier(); torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(10):
        run(train)
    t1.record(); torch.cuda.synchronize(); dist.barrier()
    res[name] = t0.elapsed_time(t1) / 10
if rank == 0:
    print(json.dumps({"world": world, **res}))
dist.destroy_process_group()
